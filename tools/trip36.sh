mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python tools/bench_configs.py f16 > gpurun_out/configs_f16.jsonl 2> gpurun_out/configs_f16.err; echo "configs f16 rc=$?" >> gpurun_out/rc.txt
timeout 900 python tools/bench_configs.py fp32 2> gpurun_out/configs_fp32.err | head -3 > gpurun_out/configs_fp32.jsonl; echo "configs fp32 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; cat gpurun_out/configs_f16.jsonl; tail -3 gpurun_out/configs_f16.err
