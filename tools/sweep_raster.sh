# A/B sweep of the tuning knobs on the bench workload (run on a B200 box)
for ro in 1 0; do for mg in 1 0; do for h in 1; do
  UGU_B200_REORDER=$ro UGU_B200_MERGE_VATTR=$mg UGU_B200_RASTER_HALVES=$h python bench.py --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('reorder $ro merge $mg halves $h', round(d['ms_per_step'],4), d['stage_us'])"
done; done; done
