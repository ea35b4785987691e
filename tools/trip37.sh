mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
for QB in 64 32 16; do
STTS_ATT_QB=$QB timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_att$QB.json 2> gpurun_out/bench_att$QB.err; echo "bench $QB rc=$?" >> gpurun_out/rc.txt
done
timeout 300 python tools/profile_step.py f16 > gpurun_out/profile_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_f16.csv python tools/profile_step.py f16 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -5 gpurun_out/pytest_gpu.log; for QB in 64 32 16; do cut -c1-200 gpurun_out/bench_att$QB.json; echo; done
