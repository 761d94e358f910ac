"""Dev: the multi-GPU table of DESIGN.md section 5 from profiles/r2_bench_n{1,2,4,8}.json (bench.py lines)."""
import json, os, sys
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles")
L = {}
for n in (1, 2, 4, 8):
    p = os.path.join(root, "r2_bench_n%d.json" % n)
    if os.path.exists(p):
        for l in open(p):
            if l.startswith("{"):
                L[n] = json.loads(l)
Ns = sorted(L)
def row(name, get, ms=None):
    cells = []
    for n in Ns:
        v = get(L[n])
        cells.append(("%.3e" % v) + ("; %.1f" % ms(L[n]) if ms else ""))
    r = get(L[Ns[-1]]) / get(L[1])
    return "| %s | %s | %.2fx |" % (name, " | ".join(cells), r)
print("| config (ray-steps/s; ms per step) | " + " | ".join("N=%d" % n for n in Ns) + " | N=%d / N=1 |" % Ns[-1])
print("|---|" + "---|" * (len(Ns) + 1))
print(row("configs[1] fisheye 2^20 rays per GPU, op6, FINAL, d_ray gathered on every GPU (WEAK)", lambda d: d["value"], lambda d: d["ms_per_step"]))
print(row("same, e2e through pinned host buffers", lambda d: d["e2e"]["value"]))
names = {"vert_64M": "configs[2] vert_heterogeneous 2^26 rays, op6, FINAL + p_x statistics (STRONG)",
         "aniso_16M_op11": "configs[3] anisotropy 2^24 rays, op11 (STRONG)", "aniso_16M_op10": "configs[3] anisotropy 2^24 rays, op10 (STRONG)",
         "shoot_4096x4096x200": "configs[4] shooting search 4096 x 4096 x 200, op6 (STRONG)",
         "interface_default": "configs[0] interface default fan (42 rays), op6, FULL s_ray + Snell errors (rank 0 only: a latency chain)"}
for k, nm in names.items():
    print(row(nm, lambda d, k=k: d["configs"][k]["value"], lambda d, k=k: d["configs"][k]["ms_per_step"]))
    if k.startswith(("vert", "aniso")):
        print(row("same, e2e through pinned host buffers", lambda d, k=k: d["configs"][k]["e2e"]["value"]))
for n in Ns:
    if n > 1:
        print("N=%d gather_check:" % n, L[n].get("gather_check"), {k: v.get("gather_check") for k, v in L[n]["configs"].items() if v.get("gather_check")})
