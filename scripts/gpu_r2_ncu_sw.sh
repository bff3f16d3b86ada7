# ncu full capture of the shallow-water kernels at 6x960x960 elements (one RHS + one JVP), after the same command ran clean
CMD="python scripts/sw_bench.py galewsky 960"
mkdir -p gpurun_out
WXB_NO_REBUILD=1 timeout 600 $CMD > gpurun_out/plain_sw.log 2>&1 && \
WXB_NO_REBUILD=1 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:sw_ -s 2 -c ${2:-2} -o gpurun_out/${1:-prof_sw} -f $CMD > gpurun_out/ncu_sw.log 2>&1
tail -2 gpurun_out/ncu_sw.log
