"""Per source line: executed instructions, stall samples and the dominant stall reasons, summed over every captured
launch of the kernels whose name contains argv[2]. Input: `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`."""
import csv, sys, collections
rows = csv.reader(open(sys.argv[1])); pat = sys.argv[2]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
fn = None; hdr = None; agg = collections.defaultdict(lambda: [0, 0, collections.Counter(), ""])
for r in rows:
    if not r: continue
    if r[0] == "Function Name": fn = r[1]; continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "File Name" or hdr is None or fn is None or pat not in fn or not r[0].isdigit(): continue
    d = dict(zip(hdr, r)); num = lambda s: int(s) if s and s.isdigit() else 0
    a = agg[int(r[0])]
    a[0] += num(d.get("Instructions Executed")); a[1] += num(d.get("# Samples")); a[3] = r[1].strip()[:90]
    for k, v in d.items():
        if k.startswith("stall_") and "Not Issued" not in k and num(v): a[2][k[6:]] += num(v)
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print(f"kernels matching {pat!r}: {ti} warp instructions, {ts} stall samples")
allst = collections.Counter()
for a in agg.values(): allst.update(a[2])
print("stall reasons overall:", ", ".join(f"{k} {100*v/max(ts,1):.0f}%" for k, v in allst.most_common(8)))
for ln, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    st = ", ".join(f"{k} {100*v/max(a[1],1):.0f}%" for k, v in a[2].most_common(3))
    print(f"{ln:5d} inst {100*a[0]/max(ti,1):5.1f}% samples {100*a[1]/max(ts,1):5.1f}% [{st}] | {a[3]}")
