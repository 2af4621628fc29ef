#!/bin/bash
# First GPU call of a round: everything that was written without a GPU, each step under its own timeout, all output into
# gpurun_out/.  Usage:  gpurun --timeout 2400 -- 'bash tools/gpu_first_call.sh'   (worst case = the sum of the step timeouts; typically ~12 min)
# (1) the regular suite; (2) the gated experimental tests (half-precision GEMM, column-blocked host path, Trrk / Syrk);
# (3) product vs the recompiled reference GPU layer; (4) blocked-host timing at the bench size; (5) the bench line.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi.txt 2>&1
timeout 240 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/pytest_gpu.log
HYD_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_local.py tests/test_gpu_gemm.py tests/test_gpu_cpp_shim.py -m gpu -q \
    -k "half or column_blocked or interdevice or basicsyrk" > gpurun_out/pytest_experimental.log 2>&1; echo "experimental rc=$?" | tee -a gpurun_out/pytest_experimental.log
HYD_TEST_EXPERIMENTAL=1 timeout 420 python -m pytest tests/test_gpu_trrk.py -m gpu -q --durations=5 > gpurun_out/pytest_trrk.log 2>&1; echo "trrk rc=$?" | tee -a gpurun_out/pytest_trrk.log
HYD_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_half.py -m gpu -q --durations=5 > gpurun_out/pytest_half.log 2>&1; echo "half rc=$?" | tee -a gpurun_out/pytest_half.log
timeout 120 tools/ref_gpu_check gpurun_out/ref_gpu_check.json 2> gpurun_out/ref_gpu_check.err; echo "ref_gpu_check rc=$?" | tee -a gpurun_out/ref_gpu_check.err
# A/B: the 128x64-tile, two-CTAs-per-SM DGEMM configuration (spill-free since the setmaxnreg split) on the same shapes
HYD_DGEMM_WN=2 timeout 120 tools/ref_gpu_check gpurun_out/ref_gpu_check_wn2.json 2> gpurun_out/ref_gpu_check_wn2.err; echo "ref_gpu_check (HYD_DGEMM_WN=2) rc=$?"
REF_CHECK_SUMMA=1 timeout 240 tools/ref_gpu_check_level1 gpurun_out/ref_gpu_check_level1.json 2> gpurun_out/ref_gpu_check_level1.err; echo "ref_gpu_check_level1 rc=$?" | tee -a gpurun_out/ref_gpu_check_level1.err
HBC_N=16384 HBC_NB=1024 HBC_TIME=1 timeout 300 python tools/host_blocked_check.py > gpurun_out/host_blocked_check.json 2> gpurun_out/host_blocked_check.err; echo "host_blocked rc=$?"
timeout 240 python tools/syrk_bench.py > gpurun_out/syrk_bench.json 2> gpurun_out/syrk_bench.err; echo "syrk_bench rc=$?"
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"

python tools/summarize_first_call.py gpurun_out > gpurun_out/SUMMARY.json 2>&1; cat gpurun_out/SUMMARY.json
