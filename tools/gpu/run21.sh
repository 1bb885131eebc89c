# r02w: 8 GPUs — bench.py under torchrun (value, e2e, multiview records with the overlapped gather), the C++
# multi-view adapter check over all 8 devices
cd $GRAFT_REPO_ROOT
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 100 --warmup 5 > gpurun_out/r02w_bench_line_8gpu.json 2> gpurun_out/r02w_bench_8gpu.err; tail -c 300 gpurun_out/r02w_bench_8gpu.err
timeout 600 tests/cpp/build/adapter_check > gpurun_out/r02w_adapter_check_8gpu.txt 2>&1; echo adapter rc=$?; tail -3 gpurun_out/r02w_adapter_check_8gpu.txt
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02w_bench_line_8gpu.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'])
print(json.dumps(d['multiview'])[:3000])
PY
