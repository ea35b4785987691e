mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_e2e_gpu.py -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
for O in -1 8 12 24 48 1000; do
STTS_OCC2=$O timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --lanes 1 --layers gpurun_out/layers_occ$O.md > gpurun_out/bench_occ$O.json 2> gpurun_out/bench_occ$O.err; echo "bench occ $O rc=$?" >> gpurun_out/rc.txt
done
cat gpurun_out/rc.txt
