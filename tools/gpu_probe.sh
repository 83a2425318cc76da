mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_overlap_input.py tests/test_generator.py tests/test_tango_api.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/probe_e2e.py 2>&1 | tee gpurun_out/probe_e2e.txt | tail -14
