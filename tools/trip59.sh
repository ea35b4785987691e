mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 4 --steps 6 --warmup 3 > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo "bench n4 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; wc -l gpurun_out/bench_n4.json; cut -c1-220 gpurun_out/bench_n4.json; tail -3 gpurun_out/bench_n4.err
