cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_visibility_tester.py -m gpu -x -q -s > gpurun_out/r02d_vt_tests.txt 2>&1; tail -30 gpurun_out/r02d_vt_tests.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02d_gpu_tests.txt 2>&1; tail -5 gpurun_out/r02d_gpu_tests.txt
