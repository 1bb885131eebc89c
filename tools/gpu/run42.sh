#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -m gpu -k "banded" > gpurun_out/r05e_banded_tests.log 2>&1; echo banded tests rc=$?; tail -3 gpurun_out/r05e_banded_tests.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r05e_tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/r05e_tests.log
for f in 1 0 1 0; do
  UGU_B200_BANDS=$f timeout 600 python bench.py --no-cpu-baseline --no-multiview 2>/dev/null | tail -1 > gpurun_out/r05e_bench_bands$f.json
  python - <<PY
import json
d=json.load(open("gpurun_out/r05e_bench_bands$f.json"))
print("bands=$f", d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["value"], d["e2e"]["pageable"]["ms_per_step"], d["e2e"]["host_link"]["ms_for_one_view_of_outputs"])
PY
done
