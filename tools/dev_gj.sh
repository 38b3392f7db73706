# dev loop for the solve kernel: correctness, timing, one ncu capture.  bash tools/dev_gj.sh <tag>
tag=${1:-dev}
python -m pytest tests/test_online_gpu.py -x -q 2>&1 | tail -3
python tools/time_sweep.py 14208 50 2>&1 | tail -8
ncu --set full --clock-control none --import-source on -k regex:k_gj_smem -s 2 -c 1 -o gpurun_out/${tag}_gj -f python tools/prof_sweep.py 14208 4 > gpurun_out/ncu_gj.log 2>&1
