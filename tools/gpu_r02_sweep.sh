# Sweep of the lane / CTA-share knobs on the final kernels (c1, driver command line otherwise).
T=${1:-r02_sweep}
mkdir -p gpurun_out
B="python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras --no-cpu-baseline --no-e2e"
for IF in 4 6 8; do for SC in 0 1 2 3; do
  timeout 200 $B --in-flight $IF --strip-ctas $SC > gpurun_out/${T}_if${IF}_sc${SC}.json 2> gpurun_out/${T}_if${IF}_sc${SC}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${T}_if${IF}_sc${SC}.json")); print("in_flight $IF strip_ctas $SC: %.1f Grays/s  ms/batch %.4f" % (d["value"]/1e3, d["config"]["ms_per_batch"]))
except Exception as e: print("in_flight $IF strip_ctas $SC failed", e)
PY
done; done
for G in 1 2; do
  timeout 200 $B --groups $G > gpurun_out/${T}_groups$G.json 2> gpurun_out/${T}_groups$G.err
  python -c "
import json; d=json.load(open('gpurun_out/${T}_groups$G.json')); print('groups $G: %.1f Grays/s' % (d['value']/1e3))"
done
