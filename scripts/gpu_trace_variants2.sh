# second A/B batch on the interface pass of csrc/euler3d.cu (fields 1): symmetric (own, facing) flux form, resident blocks
mkdir -p gpurun_out
{
for defs in "-DWXB_SYM=0" "-DWXB_SYM=1" "-DWXB_MINB4=12" "-DWXB_SYM=1 -DWXB_MINB4=12" "-DWXB_EB4=2 -DWXB_MINB4=5" "-DWXB_SYM=1 -DWXB_MINB4=8"; do
  echo "#### $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || exit 1
  bash scripts/gpu_exp.sh 0
  timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -2
done
} 2>&1 | tee gpurun_out/exp_trace_variants2.log
