# usage: bash tools/gpu_exp.sh  -- selftest + GPU tests + probe with and without programmatic dependent launch
mkdir -p gpurun_out
timeout 300 magic_b200/csrc/build/selftest_gemm > gpurun_out/selftest.log 2>&1; echo "selftest rc=$?"; grep -i -E "fail|error|PASSED" gpurun_out/selftest.log | head -20
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for pdl in 1 0 1 0; do echo "== MAGIC_PDL=$pdl"; MAGIC_PDL=$pdl timeout 300 python tests/perf_probe.py 8d 4096 bf16 2>&1 | grep -E "8d B=|sum of"; done
for e in 8c 9a 11a; do echo "== $e"; for pdl in 1 0; do MAGIC_PDL=$pdl timeout 300 python tests/perf_probe.py $e 0 bf16 2>&1 | grep -E "B=.*ms/step"; done; done
