set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands.py tests/test_visibility.py tests/test_multiview.py -m gpu -x -q > gpurun_out/r02a_tests.txt 2>&1; tail -5 gpurun_out/r02a_tests.txt
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; tail -c 600 gpurun_out/r02a_bench.json
timeout 600 python tools/bench_configs.py c1 c3 c5 c4 --no-cpu > gpurun_out/r02a_configs.jsonl 2> gpurun_out/r02a_configs.err; cat gpurun_out/r02a_configs.jsonl
UGU_B200_BATCH=1 timeout 600 python tools/bench_configs.py c1 c3 c5 --no-cpu > gpurun_out/r02a_configs_batch1.jsonl 2>&1
for b in 2 4 8; do UGU_B200_BATCH=$b timeout 300 python tools/bench_configs.py c3 c5 --no-cpu > gpurun_out/r02a_configs_batch$b.jsonl 2>&1; done
