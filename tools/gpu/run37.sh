cd $GRAFT_REPO_ROOT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/r03z_bench_line_2gpu.json 2> gpurun_out/r03z_bench_2gpu.err; echo rc=$?
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r03z_ref_2gpu.json 2>> gpurun_out/r03z_bench_2gpu.err; echo rc=$?
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r03z_bench_line_2gpu.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'])
mv=d['multiview']; print(mv['config5']['render']['ms'], mv['config5']['render_gather_to_rank0']['ms'], mv['config5']['render_fold_allreduce']['ms'], mv['config3']['render']['ms'], mv['config3']['render_gather_to_rank0']['ms'])
print(open('gpurun_out/r03z_ref_2gpu.json').read()[:200])
PY
python -c "import __graft_entry__ as g; g.smoke()"
