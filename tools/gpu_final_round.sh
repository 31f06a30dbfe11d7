# Last GPU call of the round: the full GPU suite and the default bench line on the final tree, then the ncu captures
# of the secondary kernels (shading pass, register large-scene kernel) and one c4 variant at N = 1.
T=${1:-r01_v7}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 300 python bench.py > gpurun_out/${T}_bench_c1.json 2> gpurun_out/${T}_bench_c1.err; echo c1 rc=$?; cut -c1-230 gpurun_out/${T}_bench_c1.json
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_shade -s 3 -c 1 -o gpurun_out/${T}_shade -f python tools/shade_time.py > gpurun_out/${T}_shade.log 2>&1; echo ncu-shade rc=$?
tail -1 gpurun_out/${T}_shade.log
timeout 300 ncu --set full --import-source on --clock-control none -k "regex:k_trace_level_large_reg$" -s 14 -c 1 -o gpurun_out/${T}_large_reg -f \
  python bench.py --workload c3 --large-kernel 2 --steps 1 --warmup 1 --profile > gpurun_out/${T}_large_reg.log 2>&1; echo ncu-reg rc=$?
timeout 200 python bench.py --workload c4 --in-flight 8 --slots 32 --no-e2e --no-cpu-baseline > gpurun_out/${T}_bench_c4_lanes8_slots32.json 2>/dev/null; echo c4-variant rc=$?; cut -c1-230 gpurun_out/${T}_bench_c4_lanes8_slots32.json
