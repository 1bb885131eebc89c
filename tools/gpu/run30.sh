cd $GRAFT_REPO_ROOT
for nf in 0 1; do
UGU_BENCH_NOFLUSH=$nf timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r03l_bench_nf$nf.json 2> gpurun_out/r03l_bench_nf$nf.err
python - $nf <<'PY'
import json,sys
nf=sys.argv[1]
d=json.loads(open(f'gpurun_out/r03l_bench_nf{nf}.json').read().strip().splitlines()[-1])
print('noflush',nf,'c2', round(d['ms_per_step']*1e3,1), d['ms_per_step_spread'], d['stage_us'])
PY
done
