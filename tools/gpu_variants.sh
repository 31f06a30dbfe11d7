# kernel-tuning sweep: variants/librtce_*.so (built with RTCE_NVCC_EXTRA, see build.py), each timed with the launch timeline
mkdir -p gpurun_out
cp raytracing_engine_b200/librtce.so /tmp/librtce_shipped.so
for v in variants/librtce_*.so; do
  cp $v raytracing_engine_b200/librtce.so
  echo "=== $v"
  python tools/launch_timeline.py 1 | awk '{printf "%s ", $0} /L7/&&/overflow/{print ""}' | sed 's/  */ /g' | python -c "
import sys,re
t=sys.stdin.read()
print(' '.join(m.group(1)+m.group(2)+':'+m.group(3) for m in re.finditer(r'(pix|strip|overflow) (L\d) [\d.]+ -> [\d.]+ \( ?([\d.]+) us\)', t)))
print(re.search(r'span [\d.]+ us', t).group(0))"
  timeout 120 python bench.py --no-e2e --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('bench', round(d['value']), round(d['ms_per_step']*1e3,1),'us')"
done
cp /tmp/librtce_shipped.so raytracing_engine_b200/librtce.so
