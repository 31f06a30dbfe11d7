# kernel-tuning sweep: variants/librtce_*.so (built with RTCE_NVCC_EXTRA, see build.py), each benched
mkdir -p gpurun_out
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
for v in variants/librtce_*.so; do
  cp $v raytracing_engine_b200/librtce.so
  echo "=== $v"
  for flags in "--in-flight 6" "--in-flight 1"; do
    timeout 120 python bench.py $flags --no-e2e --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('bench $flags', round(d['value']), round(d['ms_per_step']*1e3,1),'us')"
  done
done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
