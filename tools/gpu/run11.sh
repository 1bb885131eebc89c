cd $GRAFT_REPO_ROOT
export UGU_B200_TWOPASS=1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-multiview"
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --kernel-name-base demangled -k regex:raster_kernel -c 12 --csv \
  --log-file gpurun_out/r02l_launches.csv $CMD > gpurun_out/r02l_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02l_launches.csv')) if len(r)>10]
for r in rows[-9:]: print(r[4][:40], r[-3], r[-1])
PY
