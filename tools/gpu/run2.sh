set -x
cd $GRAFT_REPO_ROOT
timeout 300 python tools/bench_configs.py c3 c5 --no-cpu > gpurun_out/r02b_configs.jsonl 2>&1; cat gpurun_out/r02b_configs.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 120 --csv --log-file gpurun_out/r02b_c3_launches.csv python tools/bench_configs.py c3 --no-cpu > gpurun_out/r02b_ncu_c3.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02b_c5_launches.csv python tools/bench_configs.py c5 --views 16 --no-cpu > gpurun_out/r02b_ncu_c5.log 2>&1
