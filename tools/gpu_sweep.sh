# bench sweep over engine options: each line of $1 is a set of bench.py flags
mkdir -p gpurun_out
while read -r flags; do
  [ -z "$flags" ] && continue
  timeout 120 python bench.py $flags --no-e2e --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$flags', round(d['value']), round(d['ms_per_step']*1e3,1),'us')"
done < "$1"
