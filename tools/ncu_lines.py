"""Per-source-line totals from `ncu --page source --csv --print-source sass,cuda` (needs -lineinfo)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = []
fpath, hdr = None, None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r) if h != "Source"}
    elif hdr and fpath and r[0].isdigit() and len(r) > 8:
        try:
            out.append((int(r[hdr["Instructions Executed"]]), int(r[hdr["# Samples"]]), int(r[hdr["Thread Instructions Executed"]]), fpath, int(r[0]), r[1].strip()[:110]))
        except (ValueError, KeyError):
            pass
tot = sum(o[0] for o in out)
tots = sum(o[1] for o in out)
print(f"total warp-inst {tot:,} samples {tots:,}")
for n, s, t, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f"{100 * n / tot:6.2f}% inst {100 * s / max(tots, 1):6.2f}% smp  thr {t / max(n, 1):5.1f}  {f}:{ln}  {src}")
