mkdir -p gpurun_out; rm -f gpurun_out/rc.txt gpurun_out/e2e_errors.jsonl
for V in "A=1" "STTS_STATS_FUSED=0" "STTS_WN_FUSED=0" "STTS_OCC2=-1" "STTS_PERSISTENT=0"; do
echo "== $V" >> gpurun_out/e2e_errors.jsonl
env $V timeout 300 python -m pytest tests/test_e2e_gpu.py -q -m gpu -x -k "long_form_60s" > gpurun_out/pytest_long.log 2>&1
done
cat gpurun_out/e2e_errors.jsonl
