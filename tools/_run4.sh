( timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "fused_hex" ) > gpurun_out/t1.log 2>&1; tail -3 gpurun_out/t1.log
( timeout 600 python -m pytest tests/test_gpu_fullsize.py -x -q -k "c2" ) > gpurun_out/t2.log 2>&1; tail -3 gpurun_out/t2.log
for v in 5 12; do for d in 0; do timeout 300 python bench.py --variant $v --steps 20 --warmup 5 --no-cpu --no-e2e --dbg $d 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dbg',$d,'var',$v,j['ms_per_step'])"; done; done
timeout 300 python bench.py --variant 13 --steps 3 --warmup 1 --no-cpu --no-e2e > gpurun_out/prof.json 2> gpurun_out/prof.err; grep "femx prof" gpurun_out/prof.err | tail -1
