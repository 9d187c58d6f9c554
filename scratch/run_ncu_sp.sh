set -x
python scratch/ncu_top3.py 0 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_float_top3 -s 2 -c 1 -o gpurun_out/r02u_top3_sp2cta python scratch/ncu_top3.py 0 > gpurun_out/ncu.log 2>&1
cat gpurun_out/plain.log
