# r03s: measurement set on the final tree (same as run20): full GPU suite, C++ adapter check (+ --time), bench line +
# reference arm + launch list + ncu --set full of the per-view kernels, secondary configs, ray-cast / gbuffer / uv / d2m tools
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r03s_gpu_tests.txt 2>&1; tail -3 gpurun_out/r03s_gpu_tests.txt
timeout 600 tests/cpp/build/adapter_check > gpurun_out/r03s_adapter_check.txt 2>&1; echo adapter rc=$?; tail -2 gpurun_out/r03s_adapter_check.txt
timeout 600 tests/cpp/build/adapter_check --time > gpurun_out/r03s_adapter_check_time.txt 2>&1; echo adapter-time rc=$?
bash tools/gpu/ncu_bench.sh r03s > gpurun_out/r03s_ncu_bench.log 2>&1; tail -4 gpurun_out/r03s_ncu_bench.log
timeout 600 python tools/bench_configs.py c1 c3 c5 c4 --no-cpu > gpurun_out/r03s_configs.jsonl 2> gpurun_out/r03s_configs.err; wc -l gpurun_out/r03s_configs.jsonl
timeout 600 python tools/bench_gbuffer.py --no-cpu > gpurun_out/r03s_gbuffer_bench.jsonl 2> gpurun_out/r03s_gbuffer.err
timeout 600 python tools/bench_uvraster.py posmap vc2tex --no-cpu > gpurun_out/r03s_uvraster_bench.jsonl 2> gpurun_out/r03s_uvraster.err
timeout 600 python tools/bench_depth2mesh.py --no-cpu > gpurun_out/r03s_depth2mesh_bench.jsonl 2> gpurun_out/r03s_depth2mesh.err
ls gpurun_out | grep r03s | wc -l
