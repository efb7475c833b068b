( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/tests_gpu.log 2>&1; tail -6 gpurun_out/tests_gpu.log
timeout 600 python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -c 300 gpurun_out/bench_1gpu.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 600 gpurun_out/bench_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_l.log 2>&1; tail -3 gpurun_out/ncu_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hex_sym6 -s 1 -c 1 -f -o gpurun_out/r01_hex_sym6 python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_f.log 2>&1; tail -2 gpurun_out/ncu_f.log
timeout 300 python bench.py --variant 13 --steps 3 --warmup 1 --no-cpu --no-e2e > gpurun_out/prof.json 2> gpurun_out/prof.err; grep "femx prof" gpurun_out/prof.err | tail -1
timeout 600 python tools/config_bench.py > gpurun_out/configs_c3_c4.log 2>&1; tail -12 gpurun_out/configs_c3_c4.log
