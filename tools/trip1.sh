mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
python __graft_entry__.py > gpurun_out/build.log 2>&1; echo "build rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_gemm_gpu.py -q -k simt > gpurun_out/gemm_simt.log 2>&1; echo "gemm_simt rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m tests.stage_runner fp32 > gpurun_out/stages_fp32.log 2>&1; echo "stages_fp32 rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_gemm_gpu.py -q -k bf16 > gpurun_out/gemm_bf16.log 2>&1; echo "gemm_bf16 rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_gemm_gpu.py -q -k tf32x3 > gpurun_out/gemm_tf32.log 2>&1; echo "gemm_tf32 rc=$?" >> gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_e2e_gpu.py -q -k "fp32 or invariance or predicted or error_path or too_short or module_api" > gpurun_out/e2e_fp32.log 2>&1; echo "e2e rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
