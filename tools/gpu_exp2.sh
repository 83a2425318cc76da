# usage: bash tools/gpu_exp2.sh  -- BCE kernel: mainloop only (4) / + TMEM reads (3) / + math (1) / full (0), real data
for d in 4 7 3 1 0; do echo "== dbg $d"; MAGIC_EPI_DBG=$d PROBE_STEPS=20 timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | grep -E "fwd_dec_out_bce"; done
