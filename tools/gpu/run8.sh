# r02i: state of the tree at session re-entry — full GPU suite, C++ adapter check, the bench line + reference arm +
# launch list + ncu --set full of the four per-view kernels (tools/gpu/ncu_bench.sh), the secondary configs
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/r02i_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_gpu_tests.txt 2>&1; tail -3 gpurun_out/r02i_gpu_tests.txt
timeout 300 tests/cpp/build/adapter_check > gpurun_out/r02i_adapter_check.txt 2>&1; echo adapter rc=$?; tail -3 gpurun_out/r02i_adapter_check.txt
bash tools/gpu/ncu_bench.sh r02i > gpurun_out/r02i_ncu_bench.log 2>&1; tail -4 gpurun_out/r02i_ncu_bench.log
timeout 600 python tools/bench_configs.py c1 c3 c5 c4 --no-cpu > gpurun_out/r02i_configs.jsonl 2> gpurun_out/r02i_configs.err; wc -l gpurun_out/r02i_configs.jsonl
