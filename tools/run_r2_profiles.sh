# Round-2 ncu evidence: launch list of 4 time steps of the sweep at one full chunk (14,208 parameters, N = 64, m = 128), one
# --set full capture of the third step's kernels, and the same for the offline chain at C4 size.  bash tools/run_r2_profiles.sh
set -x
mkdir -p gpurun_out
python tools/prof_sweep.py 14208 4 > gpurun_out/prof_sweep.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_sweep_launches.csv python tools/prof_sweep.py 14208 4 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_dgemm|k_gj_smem|k_prestep' -s 6 -c 3 -o gpurun_out/r2_sweep_full -f python tools/prof_sweep.py 14208 4 > gpurun_out/ncu_f.log 2>&1
python tools/prof_offline.py > gpurun_out/prof_offline.log 2>&1 || exit 1
PROF_NO_WARMUP=1 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_offline_launches.csv python tools/prof_offline.py > gpurun_out/ncu_off_l.log 2>&1
PROF_NO_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:'k_small_gram_eig|k_small_refine|k_deim_greedy' -c 3 -o gpurun_out/r2_offline_full -f python tools/prof_treewalk.py > gpurun_out/ncu_off_f.log 2>&1
ls -la gpurun_out | tail -8
