# r02v: measurement set on the current tree: full GPU suite, bench line + reference arm + launch list + ncu --set full
# of the per-view kernels (tools/gpu/ncu_bench.sh), secondary configs with clocks
cd $GRAFT_REPO_ROOT
git_rev=$(cat gpurun_out/.rev 2>/dev/null)
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02v_gpu_tests.txt 2>&1; tail -3 gpurun_out/r02v_gpu_tests.txt
bash tools/gpu/ncu_bench.sh r02v > gpurun_out/r02v_ncu_bench.log 2>&1; tail -4 gpurun_out/r02v_ncu_bench.log
timeout 600 python tools/bench_configs.py c1 c3 c5 c4 --no-cpu > gpurun_out/r02v_configs.jsonl 2> gpurun_out/r02v_configs.err; wc -l gpurun_out/r02v_configs.jsonl
