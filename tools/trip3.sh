mkdir -p gpurun_out; rm -f gpurun_out/rc.txt gpurun_out/e2e_errors.jsonl
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
for p in fp32 tf32x3 bf16; do timeout 600 python -m tests.stage_runner $p > gpurun_out/stages_$p.log 2>&1; echo "stages_$p rc=$?" >> gpurun_out/rc.txt; done
timeout 900 python -m pytest tests/test_e2e_gpu.py -q > gpurun_out/e2e.log 2>&1; echo "e2e rc=$?" >> gpurun_out/rc.txt
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
for p in bf16 tf32x3 fp32; do timeout 900 python bench.py --steps 5 --warmup 3 --precision $p --no-cpu-baseline > gpurun_out/bench_$p.json 2> gpurun_out/bench_$p.err; echo "bench_$p rc=$?" >> gpurun_out/rc.txt; done
cat gpurun_out/rc.txt
