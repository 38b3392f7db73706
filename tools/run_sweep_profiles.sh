# ncu evidence for the online sweep at one full chunk (14,208 parameters, N = 64, m = 128): launch list of 4 steps, then one
# --set full capture of the third step's kernels.  Run on the GPU box: bash tools/run_sweep_profiles.sh <tag>
set -x
tag=${1:-r1h}
mkdir -p gpurun_out
python tools/prof_sweep.py 14208 4 > gpurun_out/prof_sweep.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_sweep_launches.csv python tools/prof_sweep.py 14208 4 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_dgemm|k_gj_smem|k_forcing|k_entries|k_tu|k_state' -s 12 -c 6 -o gpurun_out/${tag}_sweep_full -f python tools/prof_sweep.py 14208 4 > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out | tail -4
