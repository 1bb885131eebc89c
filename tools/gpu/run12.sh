cd $GRAFT_REPO_ROOT
bash tools/gpu/ncu_full.sh r03u_c3_binned 'binned_kernel' 3 python tools/bench_configs.py c3 --no-cpu
bash tools/gpu/ncu_full.sh r03u_c3_resolve 'resolve_kernel' 3 python tools/bench_configs.py c3 --no-cpu
