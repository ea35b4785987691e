mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_wn1.md > gpurun_out/bench_wn1.json 2> gpurun_out/bench_wn1.err; echo "bench rc=$?" >> gpurun_out/rc.txt
STTS_WN_FUSED=0 timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_wn0.md > gpurun_out/bench_wn0.json 2> gpurun_out/bench_wn0.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -15 gpurun_out/pytest_gpu.log; cut -c1-400 gpurun_out/bench_wn1.json; echo; cut -c1-400 gpurun_out/bench_wn0.json; tail -3 gpurun_out/bench_wn1.err
