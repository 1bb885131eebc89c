#!/bin/bash
cd /root/repo
timeout 1000 python -m pytest tests -x -q -m gpu > gpurun_out/check_tests.log 2>&1; echo tests rc=$?; tail -2 gpurun_out/check_tests.log
timeout 600 tests/cpp/build/adapter_check > gpurun_out/check_adapter.log 2>&1; echo adapter rc=$?; tail -1 gpurun_out/check_adapter.log
python bench.py > gpurun_out/check_bench_line.json 2>gpurun_out/check_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/check_bench_line.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['frac'], d['stage_us'])"
