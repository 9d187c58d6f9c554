set -x
timeout 900 python -m pytest tests/test_pose_gpu.py -x -q > gpurun_out/pose.log 2>&1; echo "pose rc=$?" >> gpurun_out/pose.log
tail -n 4 gpurun_out/pose.log
timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --steps 20 > gpurun_out/sp_bench_stereo.json 2>/dev/null; echo "bench rc=$?"
