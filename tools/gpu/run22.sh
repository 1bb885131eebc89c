cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_uvraster.py tests/test_parity_gpu.py -m gpu -x -q -k "uv or close_up or soup" > gpurun_out/r02x_tests.txt 2>&1; tail -3 gpurun_out/r02x_tests.txt
timeout 600 python tools/bench_uvraster.py posmap vc2tex --no-cpu > gpurun_out/r02x_uvraster.jsonl 2> gpurun_out/r02x_uvraster.err; python - <<'PY'
import json
for l in open('gpurun_out/r02x_uvraster.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print({k:v for k,v in d.items() if not isinstance(v,(dict,list))})
PY
