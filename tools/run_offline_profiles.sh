set -x
cd /root/repo
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r1d_bench_1gpu.json 2> gpurun_out/r1d_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r1d_bench.err
timeout 300 python tools/prof_offline.py > gpurun_out/prof_offline.log 2>&1; echo "prof rc=$?"; tail -3 gpurun_out/prof_offline.log
PROF_NO_WARMUP=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1d_offline_launches.csv python tools/prof_offline.py > gpurun_out/ncu_off_l.log 2>&1; echo "ncu launches rc=$?"
PROF_NO_WARMUP=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_deim_greedy -c 1 -o gpurun_out/r1d_greedy_full -f python tools/prof_offline.py > gpurun_out/ncu_off_g.log 2>&1; echo "ncu greedy rc=$?"
PROF_NO_WARMUP=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_dgemm -c 1 -o gpurun_out/r1d_gram_full -f python tools/prof_offline.py > gpurun_out/ncu_off_m.log 2>&1; echo "ncu gram rc=$?"
ls -la gpurun_out | tail -8
