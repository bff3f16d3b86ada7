"""Per-CUDA-source-line instruction and stall-sample shares of one kernel in an .ncu-rep (needs -lineinfo).
    python scripts/ncu_lines.py rep.ncu-rep regex:euler_main [top]"""
import collections, csv, io, subprocess, sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
first = True
inst, samp, text = collections.Counter(), collections.Counter(), {}
cur = None
ie = isamp = None
for r in rows:
    if len(r) > 7 and r[0] == "Line No" and "Instructions Executed" in r:
        if ie is not None:
            break  # only the first launch in the report
        ie, isamp = r.index("Instructions Executed"), r.index("# Samples")
        continue
    if ie is None or len(r) <= ie:
        continue
    if r[0]:
        cur = int(r[0]); text[cur] = r[1]
        continue
    if r[ie].isdigit():
        inst[cur] += int(r[ie]); samp[cur] += int(r[isamp] or 0)
tot, tots = sum(inst.values()), sum(samp.values())
print(f"total warp-instr {tot}, samples {tots}")
for ln, n in inst.most_common(top):
    print(f"{100*n/tot:5.1f}% inst {100*samp[ln]/max(tots,1):5.1f}% stall  L{ln}: {text[ln].strip()[:120]}")
