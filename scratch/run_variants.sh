for v in c1 c2n1 c2s6 c2s8; do echo "== $v"; FORESTSLAM_B200_LIB=$PWD/forest_slam_b200/variants/lib_$v.so timeout 120 python scratch/ncu_top3.py 0 1 2 3 2>&1 | tail -n 4; done
echo "== default"; timeout 120 python scratch/ncu_top3.py 0 1 2 3 2>&1 | tail -n 4
