#!/bin/bash
# fused projection: tests, adapter check, A/B bench
cd /root/repo
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r05a_tests.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/r05a_tests.log
timeout 600 tests/cpp/build/adapter_check > gpurun_out/r05a_adapter.log 2>&1; echo adapter rc=$?; tail -3 gpurun_out/r05a_adapter.log
for f in 1 0 1 0; do
  UGU_B200_FUSED=$f timeout 600 python bench.py 2>/dev/null | tail -1 > gpurun_out/r05a_bench_fused$f.json
  python - <<PY
import json
d=json.load(open("gpurun_out/r05a_bench_fused$f.json"))
print("fused=$f", d["ms_per_step"], d["value"], d.get("stage_us"), d["e2e"]["value"])
PY
done
for f in 1 0; do
  UGU_B200_FUSED=$f timeout 900 python tools/bench_configs.py c1 c3 c5 c4 --no-cpu > gpurun_out/r05a_configs_fused$f.jsonl 2>gpurun_out/r05a_configs_fused$f.err; echo configs fused=$f rc=$?
done
