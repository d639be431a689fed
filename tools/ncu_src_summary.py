"""Summarise `ncu -i X.ncu-rep --page source --csv` output: stall-reason totals, opcode histogram, hottest SASS lines.
usage: ncu -i rep --page source --csv --launch-skip 0 --launch-count 1 > /tmp/src.csv; python tools/ncu_src_summary.py /tmp/src.csv [ntop]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
H = rows[hi]
ix = {h: i for i, h in enumerate(H)}
data = [r for r in rows[hi + 1:] if len(r) >= len(H) - 2 and r[0].startswith("0x")]
I = lambda r, h: int(float(r[ix[h]] or 0))
tot_inst = sum(I(r, "Instructions Executed") for r in data)
tot_samp = sum(I(r, "# Samples") for r in data)
print("kernel:", rows[0][1][:100] if rows[0] else "")
print("total warp-inst", tot_inst, "samples", tot_samp, "sass lines", len(data))
stalls = [h for h in H if h.startswith("stall_") and "Not Issued" not in h]
st = {h: sum(I(r, h) for r in data) for h in stalls}
for h, v in sorted(st.items(), key=lambda kv: -kv[1]):
    if v:
        print("  %-25s %7d %5.1f%%" % (h, v, v / max(tot_samp, 1) * 100))
op = collections.Counter()
for r in data:
    s = r[ix["Source"]].split()
    o = s[1] if s[0].startswith("@") else s[0]
    op[o.split(".")[0]] += I(r, "Instructions Executed")
print("opcodes:")
for o, c in op.most_common(25):
    print("  %-12s %9d %5.1f%%" % (o, c, c / tot_inst * 100))
print("hottest lines by samples:")
for r in sorted(data, key=lambda r: -I(r, "# Samples"))[:ntop]:
    top = sorted(((I(r, h), h) for h in stalls), reverse=True)[:2]
    print("  %5d smp %8d exec  %-60s %s" % (I(r, "# Samples"), I(r, "Instructions Executed"), r[ix["Source"]].strip()[:60],
                                          ", ".join("%s=%d" % (h[6:], v) for v, h in top if v)))
if len(sys.argv) > 3:   # full per-line listing (address order): exec count, samples, instruction
    with open(sys.argv[3], "w") as f:
        for r in data:
            f.write("%s %9d %6d  %s\n" % (r[0][-5:], I(r, "Instructions Executed"), I(r, "# Samples"), r[ix["Source"]].strip()[:90]))
