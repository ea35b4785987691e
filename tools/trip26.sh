mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python -m pytest tests/test_gemm_gpu.py -q -m gpu -x > gpurun_out/pytest_gemm.log 2>&1; echo "pytest_gemm rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_p1.md > gpurun_out/bench_p1.json 2> gpurun_out/bench_p1.err; echo "bench rc=$?" >> gpurun_out/rc.txt
STTS_PERSISTENT=0 timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_p0.md > gpurun_out/bench_p0.json 2> gpurun_out/bench_p0.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --lanes 1 > gpurun_out/bench_p1_l1.json 2> gpurun_out/bench_p1_l1.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
