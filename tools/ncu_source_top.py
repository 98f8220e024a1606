"""Top SASS instructions by warp-stall samples from an ncu report (source page), per kernel launch id.
  python tools/ncu_source_top.py report.ncu-rep [n_top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
blocks, cur = [], []
for line in raw.splitlines():
    if line.startswith('"Kernel Name"') or line.startswith("Kernel Name"):
        if cur:
            blocks.append(cur)
        cur = [line]
    elif cur is not None:
        cur.append(line)
if cur:
    blocks.append(cur)
if not blocks:
    print(raw[:3000])
for b in blocks[:4]:
    rows = list(csv.reader(io.StringIO("\n".join(b))))
    hdr_i = next((i for i, r in enumerate(rows) if any("Sampling" in c for c in r)), None)
    if hdr_i is None:
        print("\n".join(b[:5]))
        continue
    hdr = rows[hdr_i]
    si = next(i for i, c in enumerate(hdr) if "Sampling" in c and "Not" not in c)
    src = next((i for i, c in enumerate(hdr) if c.strip() in ("Source", "SASS")), 1)
    print("== ", rows[0][:4], " columns:", hdr[:12])
    body = [r for r in rows[hdr_i + 1:] if len(r) > si]
    def num(x):
        try:
            return float(x.replace(",", ""))
        except ValueError:
            return 0.0
    tot = sum(num(r[si]) for r in body) or 1.0
    body.sort(key=lambda r: -num(r[si]))
    for r in body[:top]:
        print(f"{num(r[si]) / tot * 100:6.2f}%  {r[src][:110]}")
