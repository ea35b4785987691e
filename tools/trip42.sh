mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
nvidia-smi -L > gpurun_out/gpus.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?" >> gpurun_out/rc.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "bench ref n2 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; cut -c1-400 gpurun_out/bench_n2.json; echo; cut -c1-300 gpurun_out/bench_ref_n2.json; tail -3 gpurun_out/bench_n2.err
