# round-2 evidence, part B (1 GPU, ncu): launch list of a short bench command, then the full capture of the Euler kernel
# (one RHS + one JVP evaluation) at S2 s = 1/2; each after the same command ran clean without ncu
export WXB_NO_REBUILD=1
ARGS="--nz 12 --nh 48 --steps 3 --warmup 3 --repeats 1 --sustain-s 0 --skip-check --skip-sweep --skip-cpu --skip-sw --skip-stepops --e2e-steps 3"
mkdir -p gpurun_out
timeout 600 python bench.py $ARGS > gpurun_out/plain_b.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py $ARGS > gpurun_out/ncu_b1.log 2>&1
tail -1 gpurun_out/ncu_b1.log
timeout 600 python bench.py $ARGS > gpurun_out/plain_b.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:euler_ -s 6 -c 2 -o gpurun_out/r2_sweep_final -f python bench.py $ARGS > gpurun_out/ncu_b2.log 2>&1
tail -1 gpurun_out/ncu_b2.log
