mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
for L in 1 2 3; do
timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --lanes $L > gpurun_out/bench_lanes$L.json 2> gpurun_out/bench_lanes$L.err; echo "bench lanes $L rc=$?" >> gpurun_out/rc.txt
done
cat gpurun_out/rc.txt
