"""Per-source-line instruction / stall-sample shares from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None; cur_fn = None; hdr = None; agg = {}
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name': cur_fn = r[1][:60]; continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr is None or not r[0].isdigit(): continue
    d = dict(zip(hdr, r))
    num = lambda s: int(s) if s.isdigit() else 0
    ie = num(d.get('Instructions Executed', '0')); samp = num(d.get('# Samples', '0'))
    if ie == 0 and samp == 0: continue
    agg.setdefault(cur_fn, []).append((cur_file, int(r[0]), r[1].strip()[:100], ie, samp, d.get('Avg. Threads Executed', '')))
for fn, l in agg.items():
    tot = sum(x[3] for x in l); ts = sum(x[4] for x in l)
    print('=====', fn, 'total inst', tot, 'samples', ts)
    for x in sorted(l, key=lambda x: -x[3])[:top]:
        print(f"{x[0][:16]:16s} {x[1]:4d} {100*x[3]/tot:5.1f}% s{100*x[4]/max(ts,1):5.1f}% thr{x[5]:>4s} | {x[2]}")
