"""Dev script: static opcode histogram of the step loop (the backward-branch region holding VOTE.ANY) of a
trace kernel in an object file.  usage: sass_loop.py <obj> <mangled-kernel-substring>"""
import collections
import re
import subprocess
import sys

obj, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for f in funcs[1:]:
    name = f.split("\n", 1)[0].strip()
    if pat not in name:
        continue
    ins = []
    for m in re.finditer(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", f):
        ins.append((int(m.group(1), 16), m.group(2).strip()))
    addr_idx = {a: k for k, (a, _) in enumerate(ins)}
    vote = [k for k, (_, t) in enumerate(ins) if "VOTE" in t]
    best = None
    for k, (a, t) in enumerate(ins):
        m = re.search(r"BRA\S*\s+(?:\S+,\s*)?`\(\.L_x_\d+\)|BRA\S*\s.*0x([0-9a-f]+)", t)
        m2 = re.search(r"0x([0-9a-f]+)", t) if "BRA" in t else None
        if m2:
            tgt = int(m2.group(1), 16)
            if tgt < a and tgt in addr_idx and any(addr_idx[tgt] <= v <= k for v in vote):
                # the step loop = the backward-branch region (holding a VOTE) with the most FP64 arithmetic
                nd = sum(1 for _, tt in ins[addr_idx[tgt]:k + 1] if re.search(r"\bD(FMA|MUL|ADD)\b", tt))
                if best is None or nd > best[2]:
                    best = (addr_idx[tgt], k, nd)
    if not best:
        print(name, "no loop found"); continue
    body = ins[best[0]:best[1] + 1]
    hist = collections.Counter()
    for _, t in body:
        op = t.split()[1] if t.startswith("@") else t.split()[0]
        hist[op.split(".")[0]] += 1
    fp64 = sum(hist[k] for k in ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX"))
    print(name[:60], "loop instrs", len(body), "FP64", fp64)
    print("  " + "  ".join("%s %d" % kv for kv in hist.most_common(24)))
