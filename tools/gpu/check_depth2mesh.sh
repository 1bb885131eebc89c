#!/bin/bash
cd /root/repo
timeout 300 python -m pytest tests/test_depth2mesh.py -x -q -m gpu > gpurun_out/d2m_tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/d2m_tests.log
UGU_B200_PDL=0 timeout 300 python -m pytest tests/test_depth2mesh.py -x -q -m gpu > gpurun_out/d2m_tests_nopdl.log 2>&1; echo tests nopdl rc=$?; tail -1 gpurun_out/d2m_tests_nopdl.log
for f in 1 0 1 0; do
  UGU_B200_D2M_FUSED=$f timeout 300 python tools/bench_depth2mesh.py --no-cpu 2>/dev/null | tail -1 > gpurun_out/d2m_bench_fused$f.json
  python -c "
import json; d=json.load(open('gpurun_out/d2m_bench_fused$f.json')); print('fused=$f', {k:v for k,v in d.items() if not isinstance(v,(dict,list,str))})"
done
