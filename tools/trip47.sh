mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python tools/bench_configs.py f16 2> gpurun_out/c3a.err | head -3 | cut -c1-250 > gpurun_out/c3_default.jsonl
STTS_OCC2=-1 timeout 300 python tools/bench_configs.py f16 2> gpurun_out/c3b.err | head -3 | cut -c1-250 > gpurun_out/c3_noocc.jsonl
STTS_STATS_FUSED=0 timeout 300 python tools/bench_configs.py f16 2> gpurun_out/c3c.err | head -3 | cut -c1-250 > gpurun_out/c3_nostats.jsonl
timeout 300 python -m pytest tests/test_stages_gpu.py tests/test_e2e_gpu.py -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; for f in default noocc nostats; do echo $f; grep C3 gpurun_out/c3_$f.jsonl; grep graph gpurun_out/c3_$f.jsonl | grep C1; done
