( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/tests_gpu.log 2>&1; tail -6 gpurun_out/tests_gpu.log
timeout 600 python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -c 400 gpurun_out/bench_1gpu.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_l.log 2>&1; tail -3 gpurun_out/ncu_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hex_sym6 -s 1 -c 1 -f -o gpurun_out/r01_hex_sym6 python bench.py --variant 12 --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/ncu_f.log 2>&1; tail -2 gpurun_out/ncu_f.log; ls -la gpurun_out/*.ncu-rep
