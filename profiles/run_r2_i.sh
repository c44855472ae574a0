#!/bin/bash
# round 2, call i: new rows (examples ports, predictive path, weight variants)
mkdir -p gpurun_out/r2i
timeout 900 python -m pytest tests/test_gpu_weight_variants.py tests/test_gpu_examples.py tests/test_predictive.py -q -x --timeout 600 > gpurun_out/r2i/pytest.log 2>&1
echo "rc=$?" >> gpurun_out/r2i/pytest.log
tail -30 gpurun_out/r2i/pytest.log
