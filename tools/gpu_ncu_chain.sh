# ncu capture of the launches of one chunk's chain (1 group, launches issued, not replayed from a graph)
# usage: gpu_ncu_chain.sh <kernel regex> <skip> <count> <out name>
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"${1:-k_march_overflow|k_trace_level_pix|k_trace_strip}" -s ${2:-24} -c ${3:-12} \
  -o gpurun_out/${4:-chain} -f python bench.py --groups 1 --graph 0 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --profile > gpurun_out/${4:-chain}.log 2>&1
tail -2 gpurun_out/${4:-chain}.log | cut -c1-200
