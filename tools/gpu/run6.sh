cd $GRAFT_REPO_ROOT
timeout 300 tests/cpp/build/adapter_check > gpurun_out/r02f_adapter_check.txt 2>&1; echo rc=$?; tail -12 gpurun_out/r02f_adapter_check.txt
timeout 600 python -m pytest tests/test_visibility_tester.py tests/test_parity_gpu.py -m gpu -x -q -s > gpurun_out/r02f_tests.txt 2>&1; tail -12 gpurun_out/r02f_tests.txt
