mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
timeout 1500 ncu --set full --clock-control none --profile-from-start off -k "regex:conv_gemm_tc|adain_act|colstats|dwconv|seg_affine|polar|ola_tanh|pcph|style_fc|transpose_|expand_tokens|rowdot|adaln2|f0n_conv|upsample4|durations_k|frame_index|seg_totals|grn_scale" -o gpurun_out/r1_stts_full python tools/profile_step.py bf16 > gpurun_out/ncu_full.log 2>&1; echo "ncu_full rc=$?" >> gpurun_out/rc.txt
ls -la gpurun_out/*.ncu-rep >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
