# shading-pass variants: parity (shade tests) then time, shipped library restored afterwards
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
for v in fm fm5 fm6; do
  cp variants/librtce_$v.so raytracing_engine_b200/librtce.so
  echo "=== $v"; timeout 100 python -m pytest tests/test_parity_gpu.py -q -x -k "shade" 2>&1 | tail -1
  timeout 50 python tools/shade_time.py 2>&1 | tail -1
done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
