cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_raytrace.py -m gpu -x -q > gpurun_out/r03i_rt_tests.txt 2>&1; tail -3 gpurun_out/r03i_rt_tests.txt
timeout 600 python tools/bench_raytrace.py --no-cpu > gpurun_out/r03i_raytrace_bench.jsonl 2> gpurun_out/r03i_raytrace_bench.err; tail -3 gpurun_out/r03i_raytrace_bench.err
python - <<'PY'
import json
for l in open('gpurun_out/r03i_raytrace_bench.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print({k:(v if not isinstance(v,(dict,list)) else json.dumps(v)[:200]) for k,v in d.items()})
PY
