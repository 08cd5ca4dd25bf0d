# ncu captures of round 2 (one GPU): the wavefront sweep on the pressure factor, the wide sweep on the velocity factor,
# the velocity SpMV; each after the same command has run plain.  Raw pages are read on the CPU box with tools/ncu_summary.py.
#   gpurun --timeout 900 -- 'bash tools/round2_profiles.sh'
mkdir -p gpurun_out
P="python tools/profile_ilu.py cylinderextruded 3 1"
$P pres > gpurun_out/p_plain_pres.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_wave_sweep -s 4 -c 2 -o gpurun_out/prof_wave_sweep_r02 $P pres > gpurun_out/p_ncu_pres.log 2>&1
$P vel > gpurun_out/p_plain_vel.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_ilu_sweep -s 4 -c 2 -o gpurun_out/prof_vel_sweep_r02 $P vel > gpurun_out/p_ncu_vel.log 2>&1
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-parity --no-kernels"
$B > gpurun_out/p_plain_bench.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_spmv_rows -s 6 -c 1 -o gpurun_out/prof_spmv_vel_r02 $B > gpurun_out/p_ncu_spmv.log 2>&1
$B > gpurun_out/p_plain_bench2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 4000 --csv --log-file gpurun_out/launches_r02.csv $B > gpurun_out/p_ncu_launches.log 2>&1
tail -n 2 gpurun_out/p_ncu_*.log
