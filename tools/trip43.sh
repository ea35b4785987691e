mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
for V in 8 2; do
STTS_OCC2=$V timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --layers gpurun_out/layers_occ$V.md > gpurun_out/bench_occ$V.json 2> gpurun_out/bench_occ$V.err; echo "bench occ$V rc=$?" >> gpurun_out/rc.txt
done
cat gpurun_out/rc.txt; for V in 8 2; do cut -c1-200 gpurun_out/bench_occ$V.json; echo; done
