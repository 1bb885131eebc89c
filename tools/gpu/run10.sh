# r02k: raster order of the faces (Morton within normal buckets) x two-pass occlusion ordering — parity, then A/B
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02k_parity.txt 2>&1; tail -5 gpurun_out/r02k_parity.txt
for fo in 0 1; do for tp in 0 1; do
export UGU_B200_FACE_ORDER=$fo UGU_B200_TWOPASS=$tp
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-multiview > gpurun_out/r02k_bench_fo${fo}_tp$tp.json 2> gpurun_out/r02k_bench_fo${fo}_tp$tp.err
timeout 600 python tools/bench_configs.py c5 c4 --no-cpu > gpurun_out/r02k_configs_fo${fo}_tp$tp.jsonl 2> gpurun_out/r02k_configs_fo${fo}_tp$tp.err
done; done
python - <<'PY'
import json
for fo in (0,1):
  for tp in (0,1):
    d=json.loads(open(f'gpurun_out/r02k_bench_fo{fo}_tp{tp}.json').read().strip().splitlines()[-1])
    print('fo',fo,'tp',tp,'c2', round(d['ms_per_step']*1e3,1), d['stage_us'])
    for l in open(f'gpurun_out/r02k_configs_fo{fo}_tp{tp}.jsonl'):
        if l.startswith('{'):
            c=json.loads(l); print('fo',fo,'tp',tp,c['config'][:8], round(c['ms_per_view']*1e3,1), c['stage_us'], 'batch', round(c['batch_ms_per_view']['median']*1e3,1), c['batch_ms_per_view']['stage_us_per_view'])
PY
