# ncu full capture of the Euler kernels at the s = 1/2 bench size (one RHS + one JVP evaluation), after the same command ran clean
ARGS="--nz 12 --nh 48 --steps 3 --warmup 3 --skip-check --skip-cpu --skip-sw --skip-stepops --e2e-steps 3"
mkdir -p gpurun_out
timeout 600 python bench.py $ARGS > gpurun_out/plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:euler_ -s ${2:-6} -c ${3:-2} -o gpurun_out/${1:-prof_full} -f python bench.py $ARGS > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
