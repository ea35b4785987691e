mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_e2e_gpu.py -q -m gpu -x -k "full_size or long_form_60s" > gpurun_out/pytest_full.log 2>&1; echo "pytest rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -25 gpurun_out/pytest_full.log; tail -2 gpurun_out/e2e_errors.jsonl
