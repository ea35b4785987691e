mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py --layers gpurun_out/layers_f16.md > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; cat gpurun_out/smoke.log; cut -c1-250 gpurun_out/bench_default.json
