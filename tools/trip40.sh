mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_sf1.md > gpurun_out/bench_sf1.json 2> gpurun_out/bench_sf1.err; echo "bench rc=$?" >> gpurun_out/rc.txt
STTS_STATS_FUSED=0 timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sf0.json 2> gpurun_out/bench_sf0.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/bench_configs.py f16 2> gpurun_out/configs_sf.err | head -2 > gpurun_out/configs_sf.jsonl
cat gpurun_out/rc.txt; tail -12 gpurun_out/pytest_gpu.log; cut -c1-200 gpurun_out/bench_sf1.json; echo; cut -c1-200 gpurun_out/bench_sf0.json; echo; cut -c1-300 gpurun_out/configs_sf.jsonl
