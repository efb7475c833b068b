( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -k "fused_hex or c2" ) 2>&1 | tail -2
timeout 300 python bench.py --variant 13 --steps 5 --warmup 2 --no-cpu --no-e2e 2> gpurun_out/prof.err | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms_per_step'])"; grep "femx prof" gpurun_out/prof.err | tail -1
timeout 300 python bench.py --variant 12 --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; print('v12',json.loads(sys.stdin.read().strip().splitlines()[-1])['ms_per_step'])"
