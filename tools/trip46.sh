mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py --layers gpurun_out/layers_f16.md > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py f16 > gpurun_out/profile_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_f16.csv python tools/profile_step.py f16 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
for L in stft16 gen.cnx0.pw1 flow3.wn1 dec.decode0.conv2; do
STTS_PROFILE_LAYER=$L timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -c 1 -o /tmp/r1d_$L python tools/profile_step.py f16 > gpurun_out/ncu_$L.log 2>&1; echo "ncu_$L rc=$?" >> gpurun_out/rc.txt
ncu -i /tmp/r1d_$L.ncu-rep --page raw --csv > gpurun_out/r1d_$L.raw.csv 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page details > gpurun_out/r1d_$L.details.txt 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page source --csv > gpurun_out/r1d_$L.source.csv 2>/dev/null
done
cp /tmp/r1d_flow3.wn1.ncu-rep gpurun_out/r1_full_flow3.wn1.ncu-rep
timeout 600 python tools/bench_configs.py f16 > gpurun_out/configs_f16.jsonl 2> gpurun_out/configs_f16.err; echo "configs rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -3 gpurun_out/pytest_gpu.log; cut -c1-300 gpurun_out/bench_default.json
