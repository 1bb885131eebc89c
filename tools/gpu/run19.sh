cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_raytrace.py tests/test_parity_gpu.py -m gpu -x -q -k "host_form or hazard or golden" > gpurun_out/r02u_tests.txt 2>&1; tail -3 gpurun_out/r02u_tests.txt
bash tools/gpu/ncu_full.sh r02u_raster_ordered 'raster_kernel' 4 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview
bash tools/gpu/ncu_full.sh r02u_resolve 'resolve_kernel' 4 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview
