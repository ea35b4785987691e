mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?" >> gpurun_out/rc.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "bench ref n2 rc=$?" >> gpurun_out/rc.txt
CUDA_VISIBLE_DEVICES=0 timeout 600 python tools/bench_configs.py f16 > gpurun_out/configs_f16.jsonl 2> gpurun_out/configs_f16.err; echo "configs rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; wc -l gpurun_out/bench_n2.json gpurun_out/bench_ref_n2.json; cut -c1-200 gpurun_out/bench_n2.json; cut -c1-230 gpurun_out/configs_f16.jsonl
