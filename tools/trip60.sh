mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cp.json 2> gpurun_out/bench_cp.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/bench_configs.py f16 2> gpurun_out/cfg.err | head -3 > gpurun_out/configs_cp.jsonl
cat gpurun_out/rc.txt; tail -3 gpurun_out/pytest_gpu.log; cut -c1-200 gpurun_out/bench_cp.json; echo; cut -c1-230 gpurun_out/configs_cp.jsonl
