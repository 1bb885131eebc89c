#!/bin/bash
cd /root/repo
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_gbuffer.py -x -q -m gpu > gpurun_out/r05f_tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/r05f_tests.log
for f in 1 0 1 0; do
  UGU_B200_MAPPED_COUNTERS=$f timeout 600 python bench.py --no-cpu-baseline --no-multiview 2>/dev/null | tail -1 > gpurun_out/r05f_bench_mapped$f.json
  python - <<PY
import json
d=json.load(open("gpurun_out/r05f_bench_mapped$f.json"))
print("mapped=$f", d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["value"], d["e2e"]["pageable"]["ms_per_step"], d["e2e"]["host_link"]["ms_for_one_view_of_outputs"])
PY
done
