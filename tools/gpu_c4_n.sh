# configs[4] bench line at N GPUs (run under gpurun --gpus N)
N=${1:-8}
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus $N --workload c4 --no-cpu-baseline > gpurun_out/scale_c4_n$N.json 2> gpurun_out/scale_c4_n$N.err
echo "c4 n=$N rc=$?"
python - gpurun_out/scale_c4_n$N.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    c = d["config"]
    print(round(d["value"] / 1e3, 1), "Grays/s", round(d["ms_per_step"], 2), "ms/step", c.get("gather_check"), c.get("bands"), "e2e", d.get("e2e"))
except Exception as e:
    print("unreadable", e)
PY
tail -3 gpurun_out/scale_c4_n$N.err
