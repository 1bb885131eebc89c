cd $GRAFT_REPO_ROOT
export UGU_B200_TWOPASS=1
bash tools/gpu/ncu_full.sh r02o_raster_ordered 'raster_kernel' 4 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview
