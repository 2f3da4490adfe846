"""Per-source-line stall samples of an ncu report, ignoring barrier waits:
python tools/ncu_nonbarrier.py report.ncu-rep [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hs = None; lines = []; fname = ''
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': fname = r[1].split('/')[-1]; continue
    if len(r) > 5 and r[0] == 'Line No': hs = r; continue
    if hs is None or len(r) < 10 or not r[0]: continue
    def f(n):
        try: return float(r[hs.index(n)] or 0)
        except Exception: return 0.0
    st = {h[6:]: f(h) for h in hs if h.startswith('stall_') and 'Not Issued' not in h}
    lines.append((fname, int(r[0]), r[1].strip()[:80], f('# Samples'), f('Instructions Executed'), st))
tot = sum(l[3] for l in lines)
nb = [(l, l[3] - l[5].get('barrier', 0)) for l in lines]
tnb = sum(x[1] for x in nb)
print('total samples', tot, 'non-barrier', tnb)
for l, v in sorted(nb, key=lambda x: -x[1])[:top]:
    st = sorted(((k, x) for k, x in l[5].items() if k != 'barrier'), key=lambda kv: -kv[1])[:3]
    print(f"{v/tnb*100:5.1f}% {l[0]}:{l[1]:<4d} {l[2][:64]:64s} {[(k, int(x)) for k, x in st]}")
