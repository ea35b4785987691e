mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
STTS_PERSISTENT=all timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_pall.md > gpurun_out/bench_pall.json 2> gpurun_out/bench_pall.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; cut -c1-300 gpurun_out/bench_pall.json; tail -3 gpurun_out/bench_pall.err
