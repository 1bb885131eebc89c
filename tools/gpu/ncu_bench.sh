# usage: tools/gpu/ncu_bench.sh <tag>   — the bench line, the ncu launch list and one `--set full` capture of the
# four per-view kernels of the SAME command, in that order (each only after the plain run exited 0)
set -x
cd $GRAFT_REPO_ROOT
tag=$1
python bench.py > gpurun_out/${tag}_bench_line.json 2> gpurun_out/${tag}_bench.err || { tail -20 gpurun_out/${tag}_bench.err; exit 1; }
cat gpurun_out/${tag}_bench_line.json
python bench.py --impl reference --steps 8 --warmup 1 > gpurun_out/${tag}_bench_line_reference_arm.json 2>> gpurun_out/${tag}_bench.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview"
$CMD > /dev/null 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 400 --csv \
  --log-file gpurun_out/${tag}_launches_bench_steps3.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k "regex:vertex_kernel_b|classify_kernel|raster_kernel|binned_kernel|resolve_kernel" -s 12 -c 8 -f -o gpurun_out/${tag}_bench $CMD \
  > gpurun_out/${tag}_ncu_full.log 2>&1
tail -3 gpurun_out/${tag}_ncu_full.log
ls -la gpurun_out/
