# One measurement round on the GPU box: tests, smoke, bench lines of every workload and of the reference arm,
# the ncu launch list of the bench command and one full capture of the dominant kernel.
# usage: gpu_profile_round.sh <tag>     (outputs under gpurun_out/<tag>_*)
T=${1:-round}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -2 gpurun_out/${T}_smoke.log
timeout 300 python bench.py > gpurun_out/${T}_bench_c1.json 2> gpurun_out/${T}_bench_c1.err; echo c1 rc=$?
timeout 300 python bench.py --in-flight 1 --no-cpu-baseline > gpurun_out/${T}_bench_c1_inflight1.json 2>/dev/null; echo c1-strict rc=$?
for w in c2 c3 c4; do
  timeout 600 python bench.py --workload $w > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err; echo $w rc=$?
done
timeout 600 python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err; echo ref rc=$?
# launch list of the same command (ncu serialises the launches and runs them cold: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 3 --warmup 3 --profile > gpurun_out/${T}_launches.log 2>&1; echo launches rc=$?
# full capture of the dominant kernel: the final-level strip launch (one lane, one group, 16 frames)
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_trace_strip -s 15 -c 1 -o gpurun_out/${T}_dominant -f \
  python bench.py --groups 1 --in-flight 1 --graph 0 --steps 2 --warmup 3 --profile > gpurun_out/${T}_dominant.log 2>&1; echo dominant rc=$?
for f in gpurun_out/${T}_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    r = d.get("roofline") or {}
    print(sys.argv[1].split("/")[-1], round(d["value"], 1), d["unit"], "ms/step", round(d["ms_per_step"], 4),
          "frac", round(r.get("frac", 0), 3), "e2e", round((d.get("e2e") or {}).get("value", 0)))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
