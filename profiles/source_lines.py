"""Instructions executed and stall samples per CUDA source line of an ncu report (needs -lineinfo + --import-source on).
usage: python profiles/source_lines.py report.ncu-rep [top_n]"""
import collections
import csv
import io
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
h = rows[hi]
ie, sm = h.index("Instructions Executed"), h.index("# Samples")
cnt, smp, text = collections.Counter(), collections.Counter(), {}
for r in rows[hi + 1:]:
    if len(r) <= ie or not r[0].strip():
        continue
    try:
        n = int(r[ie])
    except ValueError:
        continue
    cnt[r[0]] += n
    smp[r[0]] += int(r[sm]) if r[sm].strip().isdigit() else 0
    text[r[0]] = r[1]
tot, stot = sum(cnt.values()), sum(smp.values())
print(f"warp instructions {tot}, samples {stot}")
for l, n in cnt.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 40):
    print(f"{n / tot * 100:5.1f}% inst {smp[l] / max(stot, 1) * 100:5.1f}% smp  L{l:>4} {text[l][:100]}")
