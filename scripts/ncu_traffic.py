"""Record the DRAM bytes of one RHS and one JVP evaluation from an `ncu --set full` report into profiles/ncu_traffic.json,
keyed by the Euler path ("sweep" / "split") and stamped with the hash of the kernel sources bench.py checks.
    python scripts/ncu_traffic.py gpurun_out/x.ncu-rep <path> <nodes> profiles/<summary>.txt"""
import csv, io, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rep, path, nodes, report = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, data = rows[0], rows[2:]
iname, ird, iwr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
units = rows[1]
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tot = {"rhs": 0.0, "jvp": 0.0}
seen = {"rhs": set(), "jvp": set()}
for r in data:
    name = r[iname]
    if "euler_" not in name:
        continue
    # euler_sweep<N, JVP>, euler_main<N, JVP>: the bool tells; euler_traces<N>: RHS and JVP launches alternate
    if "euler_traces" in name:
        key = "rhs" if "traces" not in seen["rhs"] else "jvp"
        tag = "traces"
    else:
        key = "jvp" if name.rstrip(">(EvalArgs)").endswith("1") or ", 1>" in name or "(bool)1" in name else "rhs"
        tag = name.split("<")[0]
    if tag in seen[key]:
        continue
    seen[key].add(tag)
    tot[key] += float(r[ird]) * scale[units[ird]] + float(r[iwr]) * scale[units[iwr]]
fn = os.path.join(ROOT, "profiles", "ncu_traffic.json")
rec = json.load(open(fn)) if os.path.exists(fn) else {}
rec[path] = {"nodes": nodes, "rhs_bytes": tot["rhs"], "jvp_bytes": tot["jvp"], "report": report, "source_hash": bench.kernel_source_hash(),
             "rhs_bytes_per_node": tot["rhs"] / nodes, "jvp_bytes_per_node": tot["jvp"] / nodes}
json.dump(rec, open(fn, "w"), indent=1)
print(json.dumps(rec[path]))
