mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
K="regex:conv_gemm_tc|adain_act|colstats|dwconv|seg_affine|polar|ola_tanh|pcph|style_fc|transpose_|expand_tokens|rowdot|adaln2|f0n_conv|upsample4|durations_k|frame_index|seg_totals|grn_scale"
timeout 1200 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section Occupancy --section LaunchStats --section SchedulerStats --clock-control none --profile-from-start off -k "$K" -o /tmp/r1_sections python tools/profile_step.py bf16 > gpurun_out/ncu_sections.log 2>&1; echo "ncu_sections rc=$?" >> gpurun_out/rc.txt
ncu -i /tmp/r1_sections.ncu-rep --page raw --csv > gpurun_out/r1_sections_raw.csv 2>/dev/null
ls -la /tmp/r1_sections.ncu-rep >> gpurun_out/rc.txt
for S in 29 40 122 131; do
timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:conv_gemm_tc --launch-skip $S --launch-count 1 -o gpurun_out/r1_gemm_$S python tools/profile_step.py bf16 > gpurun_out/ncu_gemm_$S.log 2>&1; echo "ncu_gemm_$S rc=$?" >> gpurun_out/rc.txt
done
ls -la gpurun_out/*.ncu-rep >> gpurun_out/rc.txt
du -sh gpurun_out >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
