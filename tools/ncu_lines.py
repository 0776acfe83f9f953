"""Aggregate warp-stall samples per CUDA source line: python tools/ncu_lines.py rep [ntop]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
cur = None; hdr = None; agg = []
for r in rows:
    if r and r[0] == 'File Path': cur = r[1]; continue
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr and r and r[0].isdigit():
        iS = hdr.index('# Samples'); iex = hdr.index('Instructions Executed')
        try: agg.append((int(r[iS]), int(r[iex]), cur.split('/')[-1], int(r[0]), r[1].strip()[:100]))
        except ValueError: pass
tot = sum(a[0] for a in agg); print('total samples', tot, 'instructions executed', sum(a[1] for a in agg))
for a in sorted(agg, key=lambda x: -x[0])[:ntop]: print(a)
