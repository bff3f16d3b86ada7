# round-2 evidence, part B2 (1 GPU, ncu): full capture of four consecutive sweep launches of the device-resident loop (RHS and JVP alternate)
# at S2 s = 1/2 (10.6 M nodes), after the same command ran clean without ncu
export WXB_NO_REBUILD=1
ARGS="--nz 12 --nh1 48 --steps 3 --warmup 3 --repeats 1 --sustain-s 0 --skip-check --skip-sweep --skip-strong --skip-cpu --skip-sw --skip-stepops --e2e-steps 3"
mkdir -p gpurun_out
timeout 600 python bench.py $ARGS > gpurun_out/plain_c.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:euler_sweep -s 7 -c 4 -o gpurun_out/r2_sweep_final -f python bench.py $ARGS > gpurun_out/ncu_c.log 2>&1
tail -c 400 gpurun_out/plain_c.log; echo; tail -c 300 gpurun_out/ncu_c.log; ls -la gpurun_out/r2_sweep_final.ncu-rep
