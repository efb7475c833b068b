python -m pytest tests/test_gpu_parity.py -x -q -k "device_plan" 2>&1 | tail -15
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
FEMX_PLAN_TIMING=1 python tools/plan_bench.py c2 2>&1 | tail -12
FEMX_PLAN_TIMING=1 python tools/plan_bench.py c3 2>&1 | tail -12
python tools/plan_bench.py c4 c5
