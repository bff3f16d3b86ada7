"""Summarise an .ncu-rep (read here, no GPU): per-kernel key metrics + SASS opcode mix + stall reasons.
    python scripts/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.txt"""
import collections, csv, io, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__sass_inst_executed_op_local_st.sum",
    "l1tex__t_sector_hit_rate.pct", "sm__sass_inst_executed_op_shared_ld.sum",
]
for r in data:
    print("====", r[hdr.index("Kernel Name")], "id", r[hdr.index("ID")])
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f"  {k:75s} {r[i]:>16s} {units[i]}")
    st = [(h, float(r[i])) for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and r[i]]
    st.sort(key=lambda x: -x[1])
    print("  stalls per issue:", ", ".join(f"{h.split('stalled_')[1].split('_per_issue')[0]}={v:.2f}" for h, v in st[:7]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = src.split('"Kernel Name",')[1:]
for b in blocks:
    rr = list(csv.reader(io.StringIO(b)))
    name = rr[0][0]
    h = rr[1]
    ia, ie, isamp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    ops, samp, tot, tots = collections.Counter(), collections.Counter(), 0, 0
    for r in rr[2:]:
        if len(r) <= max(ia, ie, isamp) or not r[ie].isdigit():
            continue
        t = r[ia].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        n, s = int(r[ie]), int(r[isamp] or 0)
        ops[op] += n; samp[op] += s; tot += n; tots += s
    print("==== SASS mix:", name[:60], "warp-instr", tot, "sass lines", len(rr) - 2)
    print("  " + "  ".join(f"{op} {100*n/tot:.1f}%({100*samp[op]/max(tots,1):.0f}%)" for op, n in ops.most_common(16)))
