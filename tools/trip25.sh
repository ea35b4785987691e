mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_a.md > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?" >> gpurun_out/rc.txt
STTS_SPLIT_BN256=0 timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_b.md > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --lanes 3 > gpurun_out/bench_c.json 2> gpurun_out/bench_c.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
