timeout 300 python bench.py --headline-only --no-cpu-baseline --steps 20 > gpurun_out/roll_base.json 2>/dev/null
FORESTSLAM_B200_LIB=$PWD/forest_slam_b200/variants/lib_roll.so timeout 300 python bench.py --headline-only --no-cpu-baseline --steps 20 > gpurun_out/roll_on.json 2>/dev/null
echo done
