mkdir -p gpurun_out; rm -f gpurun_out/rc.txt gpurun_out/e2e_errors.jsonl
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
timeout 900 python -m tests.stage_runner fp32 c1 > gpurun_out/stages_c1_fp32.log 2>&1; echo "stages_c1 rc=$?" >> gpurun_out/rc.txt
timeout 900 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
