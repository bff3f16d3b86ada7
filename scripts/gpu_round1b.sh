# re-run of the bench + launch list part of the evidence run
set -x
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo rc=$?; tail -5 gpurun_out/r1_bench.err; cat gpurun_out/r1_bench.json
ARGS="--steps 3 --warmup 3 --skip-check --skip-cpu --e2e-steps 3"
timeout 600 python bench.py $ARGS > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r1_launches.csv python bench.py $ARGS > gpurun_out/ncu1.log 2>&1
tail -n 3 gpurun_out/ncu1.log
