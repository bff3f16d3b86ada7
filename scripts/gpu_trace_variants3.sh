# third A/B batch: resident blocks per SM (register budget) of euler_main<4> with the symmetric flux form
mkdir -p gpurun_out
{
for defs in "-DWXB_SYM=1 -DWXB_MINB4=8" "-DWXB_SYM=0 -DWXB_MINB4=8" "-DWXB_SYM=1 -DWXB_MINB4=9" "-DWXB_SYM=1 -DWXB_MINB4=7" "-DWXB_SYM=1 -DWXB_MINB4=6" "-DWXB_SYM=1 -DWXB_EB4=2 -DWXB_MINB4=4"; do
  echo "#### $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || exit 1
  bash scripts/gpu_exp.sh 0
done
} 2>&1 | tee gpurun_out/exp_trace_variants3.log
