set -x
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-kernels"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_r01.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ilu_sweep -s 300 -c 2 -o gpurun_out/prof_sweep_r01 $CMD > gpurun_out/ncu_sweep.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_spmv -s 40 -c 2 -o gpurun_out/prof_spmv_r01 $CMD > gpurun_out/ncu_spmv.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_assemble -s 3 -c 1 -o gpurun_out/prof_asm_r01 $CMD > gpurun_out/ncu_asm.log 2>&1
tail -3 gpurun_out/ncu_*.log; ls -la gpurun_out
