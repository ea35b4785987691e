mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_tail.md > gpurun_out/bench_tail.json 2> gpurun_out/bench_tail.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -12 gpurun_out/pytest_gpu.log; cut -c1-200 gpurun_out/bench_tail.json; tail -3 gpurun_out/bench_tail.err
