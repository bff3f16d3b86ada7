# A/B of the trace / interface formulation knobs of csrc/euler3d.cu (WXB_TRACE_FIELDS, WXB_FOLD_LOGP, WXB_TRACE_MAP,
# WXB_TR_MINB): the full GPU suite on the tree's default build, then every variant built on the box (same image: nvcc
# is there) and timed with the bench's device step
mkdir -p gpurun_out
{
echo "#### tree default"
timeout 900 python -m pytest tests -m gpu -q 2>&1 > gpurun_out/pytest_full.log; tail -25 gpurun_out/pytest_full.log
bash scripts/gpu_exp.sh 0
for defs in "-DWXB_TRACE_FIELDS=0 -DWXB_FOLD_LOGP=0 -DWXB_TRACE_MAP=0" "-DWXB_TRACE_FIELDS=0 -DWXB_TRACE_MAP=0" "-DWXB_TRACE_FIELDS=2 -DWXB_TRACE_MAP=0" \
            "-DWXB_TRACE_FIELDS=1 -DWXB_TRACE_MAP=0" "-DWXB_TRACE_FIELDS=1 -DWXB_TRACE_MAP=0 -DWXB_TR_MINB=20" "-DWXB_TRACE_FIELDS=2 -DWXB_TRACE_MAP=1"; do
  echo "#### $defs"
  WXB_NVCC_DEFS="$defs" python wxfactory_b200/build.py --force > /dev/null || exit 1
  bash scripts/gpu_exp.sh 0
  timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -2
done
} 2>&1 | tee gpurun_out/exp_trace_variants.log
