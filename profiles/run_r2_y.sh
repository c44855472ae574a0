#!/bin/bash
# round 2, call y: weight dimensions B = 3, 4 (new template instantiations): SBM kernel + model parity, then the SBM bench (no regression at B = 2)
mkdir -p gpurun_out/r2y
timeout 1500 python -m pytest tests/test_gpu_sbm.py tests/test_gpu_model_sbm.py tests/test_gpu_newapi.py tests/test_gpu_mf.py tests/test_gpu_weight_variants.py -q -x -m gpu --timeout 900 > gpurun_out/r2y/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2y/pytest.log; tail -12 gpurun_out/r2y/pytest.log
timeout 600 python bench.py --workload sbm --no-cpu-baseline --no-extras --steps 10 > gpurun_out/r2y/bench_sbm.json 2> gpurun_out/r2y/bench_sbm.err; cut -c1-200 gpurun_out/r2y/bench_sbm.json
