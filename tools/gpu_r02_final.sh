# Final single-GPU evidence of the committed tree: tests, smoke, the bench line and the reference arm (driver command
# lines), then the ncu passes of the dominant kernel (launch list + --set full), each after its command ran clean.
T=${1:-r02_final}
mkdir -p gpurun_out
bash tools/gpu_r02_check.sh $T
P="python bench.py --gpus 1 --steps 2 --warmup 1 --profile"
timeout 300 $P > gpurun_out/${T}_profile_plain.json 2> gpurun_out/${T}_profile_plain.err; echo plain rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv $P > gpurun_out/${T}_launches.log 2>&1; echo ncu-list rc=$?
Q="python bench.py --gpus 1 --steps 12 --warmup 3 --profile --groups 1 --in-flight 1 --graph 0"
timeout 300 $Q > gpurun_out/${T}_profile_strict.json 2> gpurun_out/${T}_profile_strict.err; echo strict rc=$?
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_trace_strip -s 43 -c 1 -o gpurun_out/${T}_strip -f $Q > gpurun_out/${T}_strip.log 2>&1; echo ncu-strip rc=$?
python tools/strip_timeline.py -1 1 > gpurun_out/${T}_strip_timeline.txt 2>&1; cat gpurun_out/${T}_strip_timeline.txt
ls -la gpurun_out/*.ncu-rep
