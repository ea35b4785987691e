mkdir -p gpurun_out; rm -f gpurun_out/rc.txt gpurun_out/e2e_errors*.jsonl
timeout 900 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
mv gpurun_out/e2e_errors.jsonl gpurun_out/e2e_errors_f16x3.jsonl
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_bf16.md > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo "bench rc=$?" >> gpurun_out/rc.txt
for V in "istft" "head" "prior" "head,istft" "head,istft,prior"; do
export STTS_ISLAND_1PASS=$V
timeout 900 python -m pytest tests/test_e2e_gpu.py -q -m gpu > "gpurun_out/pytest_1pass_$V.log" 2>&1; echo "pytest_1pass $V rc=$?" >> gpurun_out/rc.txt
mv gpurun_out/e2e_errors.jsonl "gpurun_out/e2e_errors_1pass_$V.jsonl"
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_bf16_1pass.md > gpurun_out/bench_bf16_1pass.json 2> gpurun_out/bench_bf16_1pass.err; echo "bench1 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
