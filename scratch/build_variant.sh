#!/bin/bash
# build forest_slam_b200/variants/lib_<name>.so with extra -D flags for float_top3.cu (tuning experiments)
# usage: scratch/build_variant.sh name "-DT3_2CTA=1 -DT3_STAGES_N=8"
set -e
cd "$(dirname "$0")/.."
name=$1; flags=$2
mkdir -p forest_slam_b200/variants /tmp/fsvar_$name
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr $flags -c forest_slam_b200/csrc/float_top3.cu -o /tmp/fsvar_$name/float_top3.o
objs=$(ls forest_slam_b200/build/*.o | grep -v float_top3.o)
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o forest_slam_b200/variants/lib_$name.so $objs /tmp/fsvar_$name/float_top3.o
echo built forest_slam_b200/variants/lib_$name.so
