mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_bf16.md > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/bench_lanes1.json 2> gpurun_out/bench_lanes1.err; echo "bench1 rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py bf16 > gpurun_out/profile_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_bf16.csv python tools/profile_step.py bf16 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
