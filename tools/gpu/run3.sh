cd $GRAFT_REPO_ROOT
bash tools/gpu/ncu_full.sh r02c_c3_binned16 'binned_kernel.*16>' 1 python tools/bench_configs.py c3 --no-cpu
bash tools/gpu/ncu_full.sh r02c_c3_resolve16 'resolve_kernel.*16>' 1 python tools/bench_configs.py c3 --no-cpu
bash tools/gpu/ncu_full.sh r02c_c5_raster16 'raster_kernel.*16>' 1 python tools/bench_configs.py c5 --views 32 --no-cpu
