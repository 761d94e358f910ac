"""Dev script: two-point shooting search + DELTA_S sweep at configs[4] scale (4096 targets x 4096 angles,
interface scenario), timing per phase."""
import sys, os, time  # usage: shoot_probe.py [C] [A=Tn] [groups...]
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import raytracing_b200 as rt

C = int(sys.argv[1]) if len(sys.argv) > 1 else 8
A = Tn = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
eng = rt.get_engine(0)
uc, m = "1", 6
z, grd = rt.build_field(uc); eng.set_field(z, grd)
c = rt.scenario_constants(uc); gamma, box = c[0], c[5:9]
divisors, options, _ = rt.sweep_candidates(uc)
sel = np.linspace(0, len(options) - 1, C).astype(int)
steps = options[sel]
msz = rt.candidate_rows(uc, steps)
angles = np.linspace(2 * np.pi / 60, np.pi / 2, A)
true = np.linspace(0.2, 1.5, Tn) + 1.3e-4
tx, ty, T_true = rt.targets_from_angles(eng, m, uc, true, steps[0])
d = lambda a: torch.as_tensor(a).cuda()
groups = [int(g) for g in sys.argv[3:]] or [0]
for G in groups * 2:
  eng.set_option("shoot_groups", G)
  for it in range(1):
      t0 = time.perf_counter()
      co = eng.shoot_coarse(m, d(angles), (-2., -2.), d(steps), d(msz), (d(tx), d(ty)), gamma, box)
      lo, hi = rt.brackets_from_index(co.index, d(angles))
      rf = eng.shoot_refine(m, lo, hi, (-2., -2.), d(steps), d(msz), (d(tx), d(ty)), gamma, box)
      torch.cuda.synchronize()
      dt = time.perf_counter() - t0
      rs = co.stats["ray_steps"] + rf.stats["ray_steps"]
      print("G=%d C=%d A=%d Tn=%d: coarse %.2f ms (%.3e steps), refine %.2f ms (%.3e steps, %d iters, %d launches), wall %.1f ms, %.3e ray-steps/s"
            % (G, C, A, Tn, co.stats["kernel_ms"], co.stats["ray_steps"], rf.stats["kernel_ms"], rf.stats["ray_steps"], rf.iterations,
               rf.stats["launches"], dt * 1e3, rs / dt), flush=True)
th = rf.theta.cpu().numpy(); res = rf.result.cpu().numpy()
print("max |theta - true| (finest):", np.abs(th[0] - true).max(), "max miss:", res[0].max(), "max|T-Ttrue| finest", np.abs(res[1][0] - T_true).max())
print("T deviation vs finest per candidate:", np.abs(res[1] - res[1][0]).max(axis=1))
