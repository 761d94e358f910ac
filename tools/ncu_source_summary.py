"""Dev: condense `ncu -i X.ncu-rep --page source --csv` (SASS view) into per-opcode executed counts per warp-step, the
share of the stall samples, and the listing of the instructions that run every step.  usage: ncu_source_summary.py CSV RAY_STEPS"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
steps = int(sys.argv[2]) / 32
name = rows[0][1] if len(rows[0]) > 1 else ""
hdr, data = rows[1], rows[2:]
for n_, r_ in enumerate(data):   # a capture of several launches: keep the first listing
    if len(r_) < len(hdr) or r_ == hdr:
        data = data[:n_]
        break
isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
print("# ncu source page (SASS) of %s" % name)
print("# condensed by tools/ncu_source_summary.py from the capture of tools/gpu_ncu_final.sh (ncu --set full --import-source on);")
print("# per warp-step = executions / (ray-steps / 32).  ncu's listing does not print operand-reuse flags; cuobjdump -sass does")
print("# (profiles/r2_sass_loop_op6_final.txt)")
h, hs, tot, totsmp = collections.Counter(), collections.Counter(), 0.0, 0.0
for r in data:
    t = r[isrc].strip()
    ex, sm = float(r[iex]), float(r[ismp] or 0)
    o = (t.split()[1] if t.startswith("@") else t.split()[0]).split(".")[0]
    h[o] += ex / steps; hs[o] += sm; tot += ex / steps; totsmp += sm
print("\n## executed warp-instructions per warp-step by opcode (total %.1f) and share of the stall samples" % tot)
for o, v in h.most_common():
    if v >= 0.05:
        print("%-10s %7.2f   %5.1f %%" % (o, v, 100 * hs[o] / totsmp))
print("FP64 arithmetic (DFMA + DMUL + DADD) per warp-step: %.1f; DSETP %.2f" % (h["DFMA"] + h["DMUL"] + h["DADD"], h["DSETP"]))
print("\n## the step loop as executed (instructions run >= 0.9 times per warp-step), with stall samples")
print("%8s %8s  %s" % ("exec/ws", "samples", "SASS"))
for r in data:
    ex = float(r[iex]) / steps
    if ex >= 0.9:
        print("%8.2f %8d  %s" % (ex, int(float(r[ismp] or 0)), r[isrc].strip()))
