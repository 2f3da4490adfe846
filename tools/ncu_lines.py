"""Per-source-line summary of an ncu report: python tools/ncu_lines.py report.ncu-rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
lines = []; fname = ""; hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No": hdr = {h: i for i, h in enumerate(r)}; hs = r; continue
    if hdr is None or len(r) < 10 or not r[0]: continue
    def f(name):
        try: return float(r[hs.index(name)] or 0)
        except Exception: return 0.0
    stalls = {h[6:]: f(h) for h in hs if h.startswith("stall_") and "Not Issued" not in h}
    lines.append((fname, int(r[0]), r[1].strip()[:100], f("# Samples"), f("Instructions Executed"), stalls))
ts = sum(l[3] for l in lines) or 1; ti = sum(l[4] for l in lines) or 1
print(f"total samples {ts:.0f}, warp-instructions {ti:.0f}")
print("== by samples")
for l in sorted(lines, key=lambda x: -x[3])[:top]:
    st = sorted(l[5].items(), key=lambda kv: -kv[1])[:3]
    print(f"{l[3]/ts*100:5.1f}% smp {l[4]/ti*100:5.1f}% ins  {l[0]}:{l[1]:<4d} {l[2]}   [{', '.join(f'{k} {v/max(l[3],1)*100:.0f}%' for k, v in st)}]")
print("== by instructions")
for l in sorted(lines, key=lambda x: -x[4])[:top]:
    print(f"{l[4]/ti*100:5.1f}% ins {l[3]/ts*100:5.1f}% smp  {l[0]}:{l[1]:<4d} {l[2]}")
