# Large-scene (objCount > 8) round on the GPU box: parity tests, c3 bench with both kernels, one ncu capture of each.
# usage: gpu_large_round.sh <tag>
T=${1:-large}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${T}_pytest_gpu.log
timeout 400 python bench.py --workload c3 --large-kernel 2 > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench_c3.err; echo c3-reg rc=$?
timeout 400 python bench.py --workload c3 --large-kernel 1 --no-cpu-baseline > gpurun_out/${T}_bench_c3_smem.json 2> gpurun_out/${T}_bench_c3_smem.err; echo c3-smem rc=$?
for lv in 1 4; do
  timeout 300 python bench.py --workload c3 --levels $lv --no-cpu-baseline --no-e2e > gpurun_out/${T}_bench_c3_levels$lv.json 2>/dev/null; echo c3-levels$lv rc=$?
done
# ncu: level 6 (960 x 544 rays) of the first traced frame, one capture per kernel version
timeout 600 ncu --set full --import-source on --clock-control none -k "regex:k_trace_level_large_reg$" -s 14 -c 1 -o gpurun_out/${T}_large_reg -f \
  python bench.py --workload c3 --large-kernel 2 --steps 1 --warmup 1 --profile > gpurun_out/${T}_large_reg.log 2>&1; echo ncu-reg rc=$?
timeout 600 ncu --set full --import-source on --clock-control none -k "regex:k_trace_level_large$" -s 14 -c 1 -o gpurun_out/${T}_large_smem -f \
  python bench.py --workload c3 --large-kernel 1 --steps 1 --warmup 1 --profile > gpurun_out/${T}_large_smem.log 2>&1; echo ncu-smem rc=$?
for f in gpurun_out/${T}_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    r = d.get("roofline") or {}
    print(sys.argv[1].split("/")[-1], round(d["value"], 2), d["unit"], "ms/step", round(d["ms_per_step"], 3),
          "frac", round(r.get("frac", 0), 3), "step frac", round(((r.get("step") or {}).get("frac") or 0), 3), "e2e", round((d.get("e2e") or {}).get("value", 0), 2))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
