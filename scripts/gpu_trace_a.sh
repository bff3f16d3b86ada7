# A/B of the trace-field formulation (WXB_TRACE_A, csrc/euler3d.cu): parity suite + margins + bench for the default
# build, then the bench for the previous formulation built on the box (same image: nvcc is there)
mkdir -p gpurun_out
{
echo "#### WXB_TRACE_A=1 (tree default)"
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
timeout 300 python scripts/gpu_errors.py 2>&1 | tail -12
bash scripts/gpu_exp.sh 0
echo "#### WXB_TRACE_A=0"
WXB_NVCC_DEFS="-DWXB_TRACE_A=0" python wxfactory_b200/build.py --force > /dev/null || exit 1
bash scripts/gpu_exp.sh 0
} 2>&1 | tee gpurun_out/exp_trace_a.log
