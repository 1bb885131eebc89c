cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_raytrace.py -m gpu -x -q > gpurun_out/r03p_rt_tests.txt 2>&1; tail -3 gpurun_out/r03p_rt_tests.txt
for tp in 0 1; do
UGU_B200_TWOPASS=$tp timeout 600 python tools/bench_raytrace.py --no-cpu > gpurun_out/r03p_raytrace_bench_tp$tp.jsonl 2> gpurun_out/r03p_raytrace_bench_tp$tp.err
python - $tp <<'PY'
import json,sys
for l in open(f'gpurun_out/r03p_raytrace_bench_tp{sys.argv[1]}.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print('ordering',sys.argv[1], d['ms_per_view'], d['roofline']['frac'])
PY
done
timeout 900 python tools/bench_raytrace.py > gpurun_out/r03p_raytrace_bench.jsonl 2> gpurun_out/r03p_raytrace_bench.err; tail -c 600 gpurun_out/r03p_raytrace_bench.jsonl
