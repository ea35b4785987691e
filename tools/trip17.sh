mkdir -p gpurun_out; rm -f gpurun_out/rc.txt gpurun_out/e2e_errors.jsonl
timeout 600 python -m tests.stage_runner fp32 > gpurun_out/stages_fp32.log 2>&1; echo "stages_fp32 rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m tests.stage_runner bf16 > gpurun_out/stages_bf16.log 2>&1; echo "stages_bf16 rc=$?" >> gpurun_out/rc.txt
timeout 900 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
