mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_stylefc.json 2> gpurun_out/bench_stylefc.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -k regex:style_fc --csv --log-file gpurun_out/stylefc.csv python tools/profile_step.py f16 > gpurun_out/ncu_stylefc.log 2>&1
cat gpurun_out/rc.txt; tail -4 gpurun_out/pytest_gpu.log; cut -c1-200 gpurun_out/bench_stylefc.json; tail -2 gpurun_out/stylefc.csv; tail -4 gpurun_out/e2e_errors.jsonl
