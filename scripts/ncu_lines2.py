"""Per source line of one kernel in an .ncu-rep: warp instructions executed and stall samples, aggregated over
inlined copies.   python scripts/ncu_lines2.py rep.ncu-rep 'euler_sweep<(int)4, (bool)0>' [top]"""
import collections, csv, io, subprocess, sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
inst, samp, text = collections.Counter(), collections.Counter(), {}
cur_file, cur_fn, hdr, active = None, None, None, False
tot_i = tot_s = 0
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1]
    elif r[0] == "Function Name":
        cur_fn = r[1]
        active = kern in cur_fn
    elif r[0] == "Line No":
        hdr = r
        iI, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
    elif active and hdr and r[0].isdigit():
        key = (cur_file.split("/")[-1], int(r[0]))
        try:
            n, s = int(r[iI] or 0), int(r[iS] or 0)
        except ValueError:
            continue
        inst[key] += n
        samp[key] += s
        text[key] = r[1].strip()[:110]
        tot_i += n
        tot_s += s
print(f"kernel {kern}: warp instructions {tot_i}, samples {tot_s}")
print("  inst%  samp%  file:line  source")
for key, n in sorted(inst.items(), key=lambda kv: -samp[kv[0]])[:top]:
    print(f"  {100*n/max(tot_i,1):5.1f}  {100*samp[key]/max(tot_s,1):5.1f}  {key[0]}:{key[1]}  {text[key]}")
if len(sys.argv) > 4:  # line ranges "name:lo-hi,..." of the main file, plus every other file by name
    rng = [(x.split(":")[0], int(x.split(":")[1].split("-")[0]), int(x.split(":")[1].split("-")[1])) for x in sys.argv[4].split(",")]
    agg_i, agg_s = collections.Counter(), collections.Counter()
    for (f, ln), n in inst.items():
        name = f
        if f == sys.argv[5]:
            name = next((nm for nm, lo, hi in rng if lo <= ln <= hi), "other")
        agg_i[name] += n
        agg_s[name] += samp[(f, ln)]
    for nm, n in agg_i.most_common():
        print(f"  {nm:20s} inst {100*n/tot_i:5.1f}%  samples {100*agg_s[nm]/tot_s:5.1f}%")
