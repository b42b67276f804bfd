"""Per-source-line instruction and stall-sample shares from an .ncu-rep (needs -lineinfo + --import-source on)."""
import csv, io, subprocess, sys, collections
raw = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = collections.OrderedDict(); cur = None; hdr = None
for r in csv.reader(io.StringIO(raw)):
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name': continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr is None: continue
    d = dict(zip(hdr, r))
    try:
        ln = int(d['Line No']); inst = int(d['Instructions Executed'] or 0); smp = int(d['# Samples'] or 0)
    except Exception:
        continue
    a = agg.setdefault((cur, ln), [r[1][:100], 0, 0]); a[1] += inst; a[2] += smp
ti = sum(a[1] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print('total inst', ti, 'samples', ts)
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print('%-20s %4d  inst %5.1f%%  smp %5.1f%%  %s' % (k[0], k[1], 100 * a[1] / ti, 100 * a[2] / max(ts, 1), a[0]))
