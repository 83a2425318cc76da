# usage: bash tools/gpu_ncu_pair.sh -- full-set captures of the plain first-layer wgrad kernel and of its EPI_ADAM twin doing the same work
mkdir -p gpurun_out
MAGIC_FUSE_ADAM=0 timeout 900 bash tools/gpu_ncu_one.sh wgrad_plain 'gemm_tcgen05_kernel<.bool.1, .bool.1, .int.0, .int.256, .bool.1, .bool.0>' 3
MAGIC_FUSE_ADAM=1 MAGIC_ADAM_DBG=24 timeout 900 bash tools/gpu_ncu_one.sh wgrad_adam_dbg24 'gemm_tcgen05_kernel<.bool.1, .bool.1, .int.6, .int.256, .bool.1, .bool.0>' 3
ls -la gpurun_out/*.ncu-rep
