mkdir -p gpurun_out
timeout 600 python tools/diag_long2.py > gpurun_out/diag_long2.log 2>&1; echo rc=$?; tail -8 gpurun_out/diag_long2.log
