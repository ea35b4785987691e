mkdir -p gpurun_out; rm -f gpurun_out/rc.txt gpurun_out/e2e_errors.jsonl
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m tests.stage_runner fp32 > gpurun_out/stages_fp32.log 2>&1; echo "stages_fp32 rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m tests.stage_runner tf32x3 c1 > gpurun_out/stages_tf32x3_c1.log 2>&1; echo "stages_tf32_c1 rc=$?" >> gpurun_out/rc.txt
timeout 900 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/time_token_rate.py > gpurun_out/tr_time.log 2>&1
cat gpurun_out/rc.txt
