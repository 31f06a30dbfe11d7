# full check of the shipped library, then the variants, then the ncu capture of the strip kernel
T=${1:-r02_v3}
bash tools/gpu_r02_check.sh $T
bash tools/gpu_r02_variants.sh $T
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_trace_strip -s 40 -c 4 -o gpurun_out/${T}_strip -f \
  python bench.py --steps 2 --warmup 3 --profile --graph 0 > gpurun_out/${T}_ncu_strip.log 2>&1; echo ncu rc=$?
ls -la gpurun_out/*.ncu-rep
