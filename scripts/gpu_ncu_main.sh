# ncu full capture of the main and traces kernels on a short run (nh=24), after the same command ran clean
ARGS="--steps 3 --warmup 3 --nh 24 --skip-check --skip-cpu --skip-sw --skip-stepops --e2e-steps 3"
mkdir -p gpurun_out
timeout 600 python bench.py $ARGS > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:euler_ -s 8 -c 3 -o gpurun_out/${1:-prof} -f python bench.py $ARGS > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
