# r02t: multi-view C-ABI (single process, N devices) — python gpu test + C++ adapter check; two GPUs
cd $GRAFT_REPO_ROOT
nvidia-smi -L
timeout 600 python -m pytest tests/test_multiview.py tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r02t_tests.txt 2>&1; tail -5 gpurun_out/r02t_tests.txt
timeout 600 tests/cpp/build/adapter_check > gpurun_out/r02t_adapter_check.txt 2>&1; echo adapter rc=$?; tail -4 gpurun_out/r02t_adapter_check.txt
