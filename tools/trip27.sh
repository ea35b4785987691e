mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
STTS_PERSISTENT=all timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --lanes 1 --layers gpurun_out/layers_pa.md > gpurun_out/bench_pa.json 2> gpurun_out/bench_pa.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --lanes 1 --layers gpurun_out/layers_pd.md > gpurun_out/bench_pd.json 2> gpurun_out/bench_pd.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
