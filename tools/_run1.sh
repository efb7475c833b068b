set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "fused_hex" ) > gpurun_out/t1.log 2>&1; tail -5 gpurun_out/t1.log
( time timeout 600 python -m pytest tests/test_gpu_fullsize.py -x -q -k "c2" ) > gpurun_out/t2.log 2>&1; tail -5 gpurun_out/t2.log
for v in 5 12; do timeout 600 python bench.py --variant $v --steps 20 --warmup 5 --no-cpu > gpurun_out/b_v$v.json 2> gpurun_out/b_v$v.err; tail -c 600 gpurun_out/b_v$v.json; done
for d in 4 16 20; do for v in 5 12; do timeout 300 python bench.py --variant $v --steps 10 --warmup 3 --no-cpu --no-e2e --dbg $d 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dbg',$d,'var',$v,j['ms_per_step'])"; done; done
