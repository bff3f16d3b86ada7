# round-2 evidence, part B1 (1 GPU, ncu): launch list of a short bench command at the headline size (S2 s = 1), after the
# same command ran clean without ncu.  Part B2 (scripts/gpu_r2_final_c.sh, its own call) is the full capture.
export WXB_NO_REBUILD=1
ARGS="--steps 3 --warmup 3 --repeats 1 --sustain-s 0 --skip-check --skip-sweep --skip-strong --skip-cpu --skip-sw --skip-stepops --e2e-steps 3"
mkdir -p gpurun_out
timeout 600 python bench.py $ARGS > gpurun_out/plain_b.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_launches.csv python bench.py $ARGS > gpurun_out/ncu_b1.log 2>&1
tail -c 600 gpurun_out/plain_b.log; echo; tail -c 300 gpurun_out/ncu_b1.log; wc -l gpurun_out/r2_launches.csv
