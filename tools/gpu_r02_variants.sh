# A/B of strip-kernel scheduling variants (variants/librtce_<name>.so built with RTCE_NVCC_EXTRA): parity, then the
# c1 bench line without extras.  The shipped library is restored afterwards.
T=${1:-r02_v2}
mkdir -p gpurun_out
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
for v in $(ls variants | grep '\.so$' | sed 's/librtce_//; s/\.so//'); do
  cp variants/librtce_$v.so raytracing_engine_b200/librtce.so
  echo "=== $v"
  timeout 300 python -m pytest tests/test_parity_gpu.py -q -x -k "golden_all_cases or far_miss_shortcut_is_exact or batch_equals" 2>&1 | tail -1
  timeout 120 python bench.py --steps 20 --warmup 5 --no-extras --no-e2e --no-cpu-baseline > gpurun_out/${T}_bench_$v.json 2> gpurun_out/${T}_bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${T}_bench_$v.json"))
    r=d["roofline"]
    print("$v", "value %.1f Grays/s" % (d["value"]/1e3), "ms/batch %.4f" % d["config"]["ms_per_batch"], "dom ms %.4f (min %.4f)" % (r["dominant_kernel"]["ms_per_launch"], r["dominant_kernel"]["ms_per_launch_min"]), "frac %.3f step %.3f" % (r["frac"], r["step"]["frac"]), d["config"]["verify"])
except Exception as e:
    print("$v failed", e)
PY
done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
