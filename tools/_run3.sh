timeout 300 python bench.py --variant 13 --steps 3 --warmup 1 --no-cpu --no-e2e > gpurun_out/prof.json 2> gpurun_out/prof.err; grep "femx prof" gpurun_out/prof.err | tail -2
