# r03n: striding ordered raster kernel (register prefetch of the next block's list entry and indices)
cd $GRAFT_REPO_ROOT
UGU_B200_RASTER_STRIDE=2 timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "two_pass or full_size" > gpurun_out/r03n_parity.txt 2>&1; tail -3 gpurun_out/r03n_parity.txt
for rs in 0 1 2 3 4 8; do
export UGU_B200_RASTER_STRIDE=$rs
timeout 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r03n_bench_rs$rs.json 2> gpurun_out/r03n_bench_rs$rs.err
python - $rs <<'PY'
import json,sys
rs=sys.argv[1]
d=json.loads(open(f'gpurun_out/r03n_bench_rs{rs}.json').read().strip().splitlines()[-1])
print('stride',rs,'c2', round(d['ms_per_step']*1e3,1), d['ms_per_step_spread']['median'], d['stage_us'])
PY
done
