#!/bin/bash
# Round 2, call L (1 GPU): tests + shooting search with / without speculated passes.
O=gpurun_out/r2l; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
grep -E "passed|failed|FAILED|Error|rc=" $O/pytest_gpu.log | head -20
python - <<'PY'
import sys, time, numpy as np, torch
sys.path.insert(0, ".")
import raytracing_b200 as rt
eng = rt.get_engine(0)
uc, m = "1", 6
z, grd = rt.build_field(uc); eng.set_field(z, grd)
c = rt.scenario_constants(uc); gamma, box = c[0], c[5:9]
_d, options, _ = rt.sweep_candidates(uc)
steps = options
msz = rt.candidate_rows(uc, steps)
angles = np.linspace(2 * np.pi / 60, np.pi / 2, 4096)
d = lambda a: torch.as_tensor(a).cuda()
for Tn in (4096, 512):            # the problem of one GPU at N=1 and at N=8 (refinement sharded by target)
    true = np.linspace(0.2, 1.5, 4096)[:Tn] + 1.3e-4
    tx, ty, _T = rt.targets_from_angles(eng, m, uc, true, steps[0])
    co = eng.shoot_coarse(m, d(angles), (-2., -2.), d(steps), d(msz), (d(tx), d(ty)), gamma, box)
    lo, hi = rt.brackets_from_index(co.index, d(angles))
    for mode in (1, 2, 1, 2):
        eng.set_option("shoot_speculate", mode)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        rf = eng.shoot_refine(m, lo, hi, (-2., -2.), d(steps), d(msz), (d(tx), d(ty)), gamma, box)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print("Tn=%d speculate=%d: refine %.2f ms kernel, %.2f ms wall, %d launches, %.3e useful ray-steps, %.3e /s"
              % (Tn, mode, rf.stats["kernel_ms"], dt * 1e3, rf.stats["launches"], rf.stats["ray_steps"], rf.stats["ray_steps"] / dt), flush=True)
PY
