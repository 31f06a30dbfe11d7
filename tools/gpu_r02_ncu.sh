# Round-2 evidence run (one GPU): extra bench lines, then the ncu passes (each only after its command ran clean without ncu).
T=${1:-r02_v9}
mkdir -p gpurun_out
B="python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras --no-cpu-baseline"
# the 8-sphere scene (MAX_OBJECTS), the author's alternative marchers, strict stream order
timeout 300 $B --workload c1x8 > gpurun_out/${T}_bench_c1x8.json 2> gpurun_out/${T}_bench_c1x8.err; echo c1x8 rc=$?; cut -c1-300 gpurun_out/${T}_bench_c1x8.json
for M in 1 2; do
  timeout 300 $B --marcher $M --no-e2e > gpurun_out/${T}_bench_c1_marcher$M.json 2> gpurun_out/${T}_bench_c1_marcher$M.err; echo marcher$M rc=$?; cut -c1-200 gpurun_out/${T}_bench_c1_marcher$M.json
done
timeout 300 $B --kernel 1 --no-e2e > gpurun_out/${T}_bench_c1_perlevel.json 2> gpurun_out/${T}_bench_c1_perlevel.err; echo perlevel rc=$?; cut -c1-200 gpurun_out/${T}_bench_c1_perlevel.json
timeout 300 $B --in-flight 1 --no-e2e > gpurun_out/${T}_bench_c1_inflight1.json 2> gpurun_out/${T}_bench_c1_inflight1.err; echo inflight1 rc=$?; cut -c1-200 gpurun_out/${T}_bench_c1_inflight1.json
# launch list of the default plan (cold-cache, serialised: shares, not absolutes)
P="python bench.py --gpus 1 --steps 2 --warmup 1 --profile"
timeout 300 $P > gpurun_out/${T}_profile_plain.json 2> gpurun_out/${T}_profile_plain.err; echo plain rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv $P > gpurun_out/${T}_launches.log 2>&1; echo ncu-list rc=$?
# the dominant kernel: final-level strip launch in strict stream order, one frame group, launches issued (no graph)
Q="python bench.py --gpus 1 --steps 12 --warmup 3 --profile --groups 1 --in-flight 1 --graph 0"
timeout 300 $Q > gpurun_out/${T}_profile_strict.json 2> gpurun_out/${T}_profile_strict.err; echo strict rc=$?
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_trace_strip -s 43 -c 1 -o gpurun_out/${T}_strip -f $Q > gpurun_out/${T}_strip.log 2>&1; echo ncu-strip rc=$?
# the shading pass
timeout 300 python tools/shade_time.py > gpurun_out/${T}_shade_time.log 2>&1; echo shade rc=$?; tail -1 gpurun_out/${T}_shade_time.log
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_shade -s 3 -c 1 -o gpurun_out/${T}_shade -f python tools/shade_time.py > gpurun_out/${T}_shade.log 2>&1; echo ncu-shade rc=$?
ls -la gpurun_out/*.ncu-rep
