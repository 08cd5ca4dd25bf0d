# First GPU call of the next round (one GPU, about 8 minutes of box time): everything that was written after the
# GPU budget of round 1 ran out, each step under its own timeout, results in gpurun_out/.
#   gpurun --timeout 700 -- 'bash tools/round2_first_call.sh'
mkdir -p gpurun_out
# 1. the whole parity suite, then the opt-in test of the collective combiner with ranks as threads
timeout 120 python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/r2_pytest.log
S35_TEST_COMBINER=1 timeout 90 python -m pytest tests/test_gpu_partition.py -m gpu -q -k combiner 2>&1 | tail -15 > gpurun_out/r2_combiner.log
# 2. rows-in-flight SpMV: bit identity and bandwidth on the bench matrix (writes gpurun_out/tune_spmv.json)
timeout 60 python tools/tune_spmv.py > gpurun_out/r2_tune_spmv.log 2>&1
# 3. wavefront sweeps (one block per sweep, pressure factor): parity tests under the mode, then timings
S35_ILU_MODE=wavefront timeout 150 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_step.py -m gpu -q 2>&1 | tail -15 > gpurun_out/r2_wavefront_pytest.log
S35_TUNE_TIMEOUT=60 timeout 260 python tools/tune_ilu.py cylinderextruded 3 1 wavefront > gpurun_out/r2_tune_wavefront.log 2>&1
# 4. the default bench line (roofline, kernels, e2e, cpu_baseline)
timeout 200 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
tail -c 600 gpurun_out/r2_pytest.log gpurun_out/r2_combiner.log gpurun_out/r2_tune_spmv.log gpurun_out/r2_wavefront_pytest.log gpurun_out/r2_tune_wavefront.log
