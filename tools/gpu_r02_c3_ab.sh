# configs[3] (10 000 spheres, 1080p, 8 levels, one frame) on every library under variants/, register kernel, 3 timed frames each
mkdir -p gpurun_out
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
for v in $(ls variants | grep '\.so$' | sed 's/librtce_//; s/\.so//'); do
  cp variants/librtce_$v.so raytracing_engine_b200/librtce.so
  for rep in 1 2; do
  timeout 300 python bench.py --gpus 1 --workload c3 --large-kernel 2 --steps 3 --warmup 1 --no-extras --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v', 'ms/frame %.1f' % d['ms_per_step'], d['clocks'])"
  done
done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
