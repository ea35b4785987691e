mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py bf16 > gpurun_out/profile_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_bf16.csv python tools/profile_step.py bf16 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
