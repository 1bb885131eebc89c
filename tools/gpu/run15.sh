cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02q_gpu_tests.txt 2>&1; tail -3 gpurun_out/r02q_gpu_tests.txt
timeout 300 tests/cpp/build/adapter_check > gpurun_out/r02q_adapter_check.txt 2>&1; echo adapter rc=$?; tail -2 gpurun_out/r02q_adapter_check.txt
