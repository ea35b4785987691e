mkdir -p gpurun_out
timeout 600 python tools/diag_long.py > gpurun_out/diag_long.log 2>&1; echo rc=$?; tail -12 gpurun_out/diag_long.log
