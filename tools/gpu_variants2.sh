# shading-pass / large-scene kernel variants (variants/librtce_*.so built with RTCE_NVCC_EXTRA): parity, then time
mkdir -p gpurun_out
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
echo "=== shipped"; timeout 60 python tools/shade_time.py 2>&1 | tail -1
for v in vote0 vote2; do
  cp variants/librtce_$v.so raytracing_engine_b200/librtce.so
  echo "=== $v"; timeout 120 python -m pytest tests/test_parity_gpu.py -q -x -k "shade" 2>&1 | tail -1
  timeout 60 python tools/shade_time.py 2>&1 | tail -1
done
cp variants/librtce_hoist.so raytracing_engine_b200/librtce.so
echo "=== hoist"; timeout 120 python -m pytest tests/test_parity_gpu.py -q -x -k "large" 2>&1 | tail -1
timeout 120 python bench.py --workload c3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c3 hoist', round(d['value'],3), round(d['ms_per_step'],1),'ms')"
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
