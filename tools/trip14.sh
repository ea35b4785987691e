mkdir -p gpurun_out
for CFG in none prior cnx dec. flow prior,cnx; do
rm -f gpurun_out/e2e_errors.jsonl
STTS_FORCE_ISLAND=$CFG timeout 600 python -m pytest tests/test_e2e_gpu.py -q -k "bf16" > gpurun_out/exp_$CFG.log 2>&1
echo "== $CFG" >> gpurun_out/exp_summary.txt; grep bf16 gpurun_out/e2e_errors.jsonl >> gpurun_out/exp_summary.txt
done
cat gpurun_out/exp_summary.txt
