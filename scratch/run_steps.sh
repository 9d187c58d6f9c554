for K in 10 30 60; do timeout 300 python bench.py --headline-only --no-cpu-baseline --no-verify --steps $K > gpurun_out/st_$K.json 2>/dev/null; done
CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 300 python bench.py --headline-only --no-cpu-baseline --no-verify --steps 60 --lanes 8 > gpurun_out/st_60_l8c32.json 2>/dev/null
timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --no-verify --steps 40 > gpurun_out/st_stereo40.json 2>/dev/null
echo done
