# configs[4] at N GPUs: the bench line, then two variants of the lane / chunk set-up (no e2e), then the c1 line
N=${1:-8}
mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
run 29600 --workload c4 --no-cpu-baseline > gpurun_out/scale_c4_n$N.json 2> gpurun_out/scale_c4_n$N.err; echo "c4 rc=$?"
run 29601 --workload c4 --no-cpu-baseline --no-e2e --in-flight 8 > gpurun_out/scale_c4_n${N}_lanes8.json 2>/dev/null; echo "c4 lanes8 rc=$?"
run 29602 --workload c4 --no-cpu-baseline --no-e2e --in-flight 8 --slots 32 > gpurun_out/scale_c4_n${N}_lanes8_slots32.json 2>/dev/null; echo "c4 lanes8 slots32 rc=$?"
run 29603 --workload c4 --no-cpu-baseline --no-e2e --slots 32 > gpurun_out/scale_c4_n${N}_slots32.json 2>/dev/null; echo "c4 slots32 rc=$?"
for f in gpurun_out/scale_c4_n${N}*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    b = d["config"].get("bands") or {}
    print(sys.argv[1].split("/")[-1], round(d["value"] / 1e3, 1), "Grays/s", round(d["ms_per_step"], 2), "ms/step", b.get("marched_texel_weight"), b.get("set_up_trials_ms_per_step"), b.get("rows_per_rank"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
done
