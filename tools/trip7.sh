mkdir -p gpurun_out; rm -f gpurun_out/rc.txt gpurun_out/e2e_errors.jsonl
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
timeout 900 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench_ref rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py bf16 > gpurun_out/profile_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --profile-from-start off -k regex:stts -o gpurun_out/r1_stts_full python tools/profile_step.py bf16 > gpurun_out/ncu_full.log 2>&1; echo "ncu_full rc=$?" >> gpurun_out/rc.txt
ls -la gpurun_out/*.ncu-rep >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
