# More than 8 lanes? (library built with -DRTCE_MAX_LANES=16 under variants/)
T=${1:-r02_lanes}
mkdir -p gpurun_out
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
cp variants/librtce_lanes16.so raytracing_engine_b200/librtce.so
B="python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras --no-cpu-baseline --no-e2e"
for IF in 8 10 12 16; do
  timeout 200 $B --in-flight $IF > gpurun_out/${T}_if${IF}.json 2> gpurun_out/${T}_if${IF}.err
  python -c "
import json; d=json.load(open('gpurun_out/${T}_if${IF}.json')); print('in_flight $IF: %.1f Grays/s  ms/batch %.4f' % (d['value']/1e3, d['config']['ms_per_batch']), d['config']['verify'])" || tail -3 gpurun_out/${T}_if${IF}.err
done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
