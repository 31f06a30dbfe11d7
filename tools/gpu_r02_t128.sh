# 128-thread strip CTAs (variants/librtce_t128.so) against the shipped 256-thread ones, crowded default path
mkdir -p gpurun_out
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
B="python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras --no-cpu-baseline --no-e2e"
run() { timeout 200 $B $2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$1 $2: %.1f Grays/s  ms/batch %.4f  alone %.4f' % (d['value']/1e3, d['config']['ms_per_batch'], r['dominant_kernel']['ms_per_launch']), d['config']['verify'])"; }
run base ""
cp variants/librtce_t128.so raytracing_engine_b200/librtce.so
timeout 300 python -m pytest tests/test_parity_gpu.py -q -x -k "golden_all_cases or far_miss_shortcut_is_exact or batch_equals" 2>&1 | tail -1
for SC in 1 2 3 4; do run t128 "--strip-ctas $SC"; done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
