cd $GRAFT_REPO_ROOT
bash tools/gpu/ncu_full.sh r03m_raster 'raster_kernel' 4 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview
