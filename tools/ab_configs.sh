# A/B helper for the secondary configs (run on a B200 box)
# usage: tools/ab_configs.sh "c1 c3" "VAR=1" "VAR=2" ...
cfgs="$1"; shift
for e in "$@"; do
  env $e python tools/bench_configs.py $cfgs --no-cpu 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if not l.startswith('{'): continue
    d=json.loads(l); print('$e', d['config'][:8], 'ms/view', round(d['ms_per_view'],4), d['stage_us'])"
done
