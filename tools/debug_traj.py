"""Dev: where do the staged (TMA) and per-lane trajectory stores differ?"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
import raytracing_b200 as rt
eng = rt.get_engine()
for uc, m, R, M, stride in (("2", 6, 1 << 16, 300, 1), ("2", 6, 1 << 16, 40, 1), ("3", 6, (1 << 16) + 10, 1500, 3)):
    eng.set_field(*rt.build_field(uc))
    c = rt.scenario_constants(uc)
    th, x0, y0 = (torch.from_numpy(a).cuda() for a in rt.make_fan(uc, R))
    rows = (M + stride - 1) // stride
    out = []
    for mode in (0, 1, 0, 1):
        eng.set_option("traj_store", mode)
        buf = torch.full((rows, 6, R), -3.0, dtype=torch.float64, device="cuda")
        r = eng.trace(m, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], full=True, out_stride=stride, out=dict(s_ray=buf))
        out.append(r.s_ray.clone())
    eng.set_option("traj_store", 0)
    fin = eng.trace(m, th, x0, y0, rt.DELTA_S, M, c[0], c[5:9], full=False)
    print("case", uc, m, R, M, stride, "per-lane repeat equal:", torch.equal(out[0], out[2]), "staged repeat equal:", torch.equal(out[1], out[3]),
          "per-lane == staged:", torch.equal(out[0], out[1]))
    li = fin.d_ray[2].long()
    if stride == 1:
        k = torch.arange(0, R, 997, device="cuda")
        for name, o in (("per-lane", out[0]), ("staged", out[1])):
            print("  ", name, "last row matches FINAL:", torch.equal(o[li[k], :, k].T, fin.last_row[:, k]))
    d = (out[0] != out[1])
    if d.any():
        idx = d.nonzero()
        print("   differing entries:", idx.shape[0], "rows", idx[:, 0].min().item(), "..", idx[:, 0].max().item(),
              "cols", idx[:, 1].unique().tolist(), "rays", idx[:, 2].min().item(), "..", idx[:, 2].max().item())
        for j in range(min(6, idx.shape[0])):
            a = tuple(idx[j].tolist())
            print("     ", a, out[0][a].item(), out[1][a].item())
        rws = idx[:, 0].unique()
        print("   rows with differences:", rws[:20].tolist(), "... count", rws.numel())
