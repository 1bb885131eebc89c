# A/B helper (run on a B200 box): prints ms/step + stage breakdown for each env setting given
# usage: tools/ab.sh "VAR=1" "VAR=0 OTHER=2" ...
for cfg in "$@"; do
  env $cfg python bench.py --steps 200 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$cfg', 'ms/step', round(d['ms_per_step'],4), 'profiled', round(d.get('ms_per_step_stage_timed_pass',0),4), d['stage_us'])"
done
