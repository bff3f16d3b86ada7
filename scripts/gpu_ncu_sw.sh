# ncu full capture of the shallow-water kernels (galewsky tiled to nh=480), after the same command ran clean
mkdir -p gpurun_out
timeout 300 python scripts/sw_bench.py galewsky 480 > gpurun_out/sw_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sw_ -s 6 -c 4 -o gpurun_out/${1:-prof_sw} -f python scripts/sw_bench.py galewsky 480 > gpurun_out/ncu_sw.log 2>&1
tail -2 gpurun_out/ncu_sw.log
