timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/gpu_tests_1gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/f32_bench1.log 2>&1; echo rc=$? >> gpurun_out/f32_bench1.log
tail -2 gpurun_out/f32_bench1.log | cut -c1-300
