mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m pytest tests/ -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py --layers gpurun_out/layers_f16.md > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py f16 > gpurun_out/profile_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_f16.csv python tools/profile_step.py f16 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
K="regex:conv_gemm_tc|wn_layer|adain_act|colstats|dwconv|seg_affine|polar|ola_tanh|pcph|style_fc|transpose_|expand_tokens|rowdot|adaln2|f0n_conv|upsample4|durations_k|frame_index|seg_totals|grn_|attention_|row_ln|embed_k|rows_to_dense|dense_to_rows|seg_broadcast"
timeout 1500 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section Occupancy --section LaunchStats --section SchedulerStats --clock-control none --profile-from-start off -k "$K" -o /tmp/r1_sections python tools/profile_step.py f16 > gpurun_out/ncu_sections.log 2>&1; echo "ncu_sections rc=$?" >> gpurun_out/rc.txt
ncu -i /tmp/r1_sections.ncu-rep --page raw --csv > gpurun_out/r1_sections_raw.csv 2>/dev/null
cat gpurun_out/rc.txt; tail -3 gpurun_out/pytest_gpu.log; cut -c1-700 gpurun_out/bench_default.json
