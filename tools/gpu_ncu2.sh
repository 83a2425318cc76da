mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k packed 2>&1 | tail -3
timeout 900 bash tools/gpu_ncu_one.sh wgrad_enc0_bits 'gemm_tcgen05_kernel<.bool.1, .bool.1, .int.0, .int.256, .bool.1>' 3
timeout 900 bash tools/gpu_ncu_one.sh fwd_enc0_bits 'gemm_tcgen05_kernel<.bool.0, .bool.1, .int.0, .int.256, .bool.1>' 3
ls -la gpurun_out/*.ncu-rep
