# A/B of strip-kernel variants on one box + ncu capture (final-level strip launch) of two of them
T=${1:-r02_v4}
mkdir -p gpurun_out
bash tools/gpu_r02_variants.sh $T
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
for v in old d2; do
  cp variants/librtce_$v.so raytracing_engine_b200/librtce.so
  timeout 600 ncu --set full --import-source on --clock-control none -k "regex:k_trace_strip<4, 0" -s 43 -c 1 -o gpurun_out/${T}_strip_$v -f \
    python bench.py --steps 2 --warmup 3 --profile --graph 0 --groups 1 --in-flight 1 > gpurun_out/${T}_ncu_strip_$v.log 2>&1; echo ncu $v rc=$?
done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
ls -la gpurun_out/*.ncu-rep
