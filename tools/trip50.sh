mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
for L in 2 4 6; do
timeout 300 python bench.py --steps 12 --warmup 4 --no-cpu-baseline --lanes $L > gpurun_out/bench_l$L.json 2> gpurun_out/bench_l$L.err; echo "bench lanes $L rc=$?" >> gpurun_out/rc.txt
done
cat gpurun_out/rc.txt; for L in 2 4 6; do python -c "import json;b=json.load(open('gpurun_out/bench_l$L.json'));print($L,b['value'],b['e2e']['value'],b['config']['workspace_bytes'])"; done
