cd $GRAFT_REPO_ROOT
export UGU_B200_TILES=0
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none --kernel-name-base demangled -k regex:"tile_kernel|binned_kernel|raster_kernel|resolve" -s 6 -c 3 --csv --log-file gpurun_out/r02s_c3_launches_tiles0.csv python tools/bench_configs.py c3 --no-cpu > gpurun_out/r02s_ncu_c3.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02s_c3_launches_tiles0.csv')) if len(r)>10]
for r in rows[1:]: print(r[4][:34], r[7], r[-3][:40], r[-1])
PY
