set -x
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/r1_gpu_tests.log; cat gpurun_out/r1_gpu_tests.log
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo rc=$?; tail -5 gpurun_out/r1_bench.err; cat gpurun_out/r1_bench.json
timeout 600 python bench.py --steps 3 --warmup 3 --nh 24 --skip-check --skip-cpu --e2e-steps 3 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 3 --warmup 3 --nh 24 --skip-check --skip-cpu --e2e-steps 3 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:euler_main -s 4 -c 2 -o gpurun_out/r1_main python bench.py --steps 3 --warmup 3 --nh 24 --skip-check --skip-cpu --e2e-steps 3 > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:euler_traces -s 4 -c 1 -o gpurun_out/r1_traces python bench.py --steps 3 --warmup 3 --nh 24 --skip-check --skip-cpu --e2e-steps 3 > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log
