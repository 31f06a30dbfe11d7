# ncu captures of the secondary kernels: the two large-scene kernels (level 6 of the 10k-sphere frame) and the shading pass
T=${1:-r01_v7}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
timeout 600 ncu --set full --import-source on --clock-control none -k "regex:k_trace_level_large_reg$" -s 14 -c 1 -o gpurun_out/${T}_large_reg -f \
  python bench.py --workload c3 --large-kernel 2 --steps 1 --warmup 1 --profile > gpurun_out/${T}_large_reg.log 2>&1; echo ncu-reg rc=$?
timeout 600 ncu --set full --import-source on --clock-control none -k "regex:k_trace_level_large$" -s 14 -c 1 -o gpurun_out/${T}_large_smem -f \
  python bench.py --workload c3 --large-kernel 1 --steps 1 --warmup 1 --profile > gpurun_out/${T}_large_smem.log 2>&1; echo ncu-smem rc=$?
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_shade -s 3 -c 1 -o gpurun_out/${T}_shade -f python tools/shade_time.py > gpurun_out/${T}_shade.log 2>&1; echo ncu-shade rc=$?
tail -1 gpurun_out/${T}_shade.log
ls -la gpurun_out/*.ncu-rep
