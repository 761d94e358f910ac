"""Batched two-point shooting search + DELTA_S sweep (BASELINE.json configs[4]).

NOT a reference function: RT_bench.py only shoots forward fans (README.md:6) and never calls the
scipy.optimize.minimize it imports (RT_bench.py:48).  The search is therefore defined here from
reference parts and is parity-unpinned:

  1. coarse  : the fan `angles` (A launch angles from the common source) is integrated by trazar's
               kernel for every DELTA_S candidate; each ray's exit point is where its last segment
               crosses the box of constants() (RT_bench.py:878), travel time interpolated alike;
  2. bracket : per (candidate, target) j* = first argmin_j |exit_j - target|^2, bracket
               [angles[j*-1], angles[j*+1]] clipped to the fan;
  3. refine  : the reference's golden(f, lo, hi) (RT_bench.py:175-199, tol = GOLD_TOL) with
               f(theta) = |exit(theta) - target|^2, every probe a full re-trace on the GPU;
  4. result  : theta* = (a+b)/2 re-traced once: miss distance, travel time, exit point, step count.

Everything numeric runs in librtb.so (rtb_shoot_coarse / rtb_shoot_refine); this module only slices
arrays and, across GPUs, reduces the coarse (miss^2, index) pairs with MIN (torch.distributed).
The tests hold a CPU restatement of the same procedure (test infrastructure, never imported here).
"""
import numpy as np

from . import constants as K
from .engine import max_size_for, _is_torch

INT64_MAX = np.iinfo(np.int64).max


def default_source(user_choice):
    """Common launch point of the scenario's fan, RT_bench.py:809-813."""
    uc = K.scenario_id(user_choice)
    pos_x = K.constants(uc)[3]
    if uc == "2":
        return float(pos_x[0]), float(pos_x[1])
    return float(np.asarray(pos_x).ravel()[0]), -2.0


def candidate_rows(user_choice, steps, divisors=None):
    """max_size per DELTA_S candidate, RT_bench.py:796-799."""
    uc = K.scenario_id(user_choice)
    steps = np.atleast_1d(np.asarray(steps, dtype=np.float64))
    if divisors is None:
        divisors = np.full(steps.shape, K.DELTA_S_DIVISOR_FISHEYE)
    return np.array([max_size_for(uc, s, d + 1) for s, d in zip(steps, np.atleast_1d(divisors))], dtype=np.int64)


def brackets_from_index(index, angles_all):
    """[angles[j-1], angles[j+1]] clipped to the fan; numpy or torch (same library as `index`)."""
    A = angles_all.shape[0]
    if _is_torch(index):
        import torch
        ang = angles_all if _is_torch(angles_all) else torch.as_tensor(np.asarray(angles_all)).to(index.device)
        return ang[torch.clamp(index - 1, min=0)], ang[torch.clamp(index + 1, max=A - 1)]
    ang = np.asarray(angles_all)
    return ang[np.maximum(index - 1, 0)], ang[np.minimum(index + 1, A - 1)]


def reduce_argmin(miss2, index, group=None):
    """Combine per-rank (miss^2, global index) of an angle-sharded coarse fan: global minimum, ties to
    the lowest index (np.argmin's first occurrence).  Two MIN all-reduces; in place on torch tensors."""
    import torch
    import torch.distributed as dist
    gmin = miss2.clone()
    dist.all_reduce(gmin, op=dist.ReduceOp.MIN, group=group)
    cand = torch.where(miss2 == gmin, index, torch.full_like(index, INT64_MAX))
    dist.all_reduce(cand, op=dist.ReduceOp.MIN, group=group)
    return gmin, cand


def targets_from_angles(engine, option, user_choice, true_angles, step, src=None, exact=False):
    """Synthetic two-point problem with a known answer: the exit points of the rays launched at
    `true_angles` (integrated with `step`) become the targets.  Returns (tx, ty, T) numpy arrays."""
    uc = K.scenario_id(user_choice)
    c = K.constants(uc)
    src = default_source(uc) if src is None else src
    th = np.ascontiguousarray(true_angles, dtype=np.float64)
    far = np.array([1e300]), np.array([1e300])  # one dummy target: only the exit points are wanted
    r = engine.shoot_coarse(option, th, src, [step], candidate_rows(uc, [step]), far, c[0], c[5:9], want_exit=True,
                            exact=exact)
    return r.exit[0, 0].copy(), r.exit[1, 0].copy(), r.exit[2, 0].copy()


def shooting_search(option, z, grd, user_choice, targets, angles, steps=None, divisors=None, src=None, engine=None,
                    group=None, device_arrays=False, exact=False):
    """Two-point search for every (DELTA_S candidate, target).

    targets = (tx, ty) arrays of Tn points; angles = coarse fan (A,); steps = DELTA_S candidates (C,)
    (default: the scenario's DELTA_S).  With a torch.distributed `group` (or an initialised default
    group and group=True) the coarse fan is sharded by angle and the refinement by target across ranks;
    every rank returns the full result.  Returns dict of (C, Tn) arrays: theta, miss, T, exit_x, exit_y,
    nsteps (int64), index (int64 coarse argmin), coarse_miss; plus iterations and stats.
    """
    from .api import get_engine
    eng = engine or get_engine()
    uc = K.scenario_id(user_choice)
    c = K.constants(uc)
    gamma, box = c[0], c[5:9]
    eng.set_field(z, grd)
    src = default_source(uc) if src is None else src
    steps = np.atleast_1d(np.asarray([K.DELTA_S] if steps is None else steps, dtype=np.float64))
    msz = candidate_rows(uc, steps, divisors)
    tx, ty = (np.ascontiguousarray(t, dtype=np.float64) for t in targets)
    angles = np.ascontiguousarray(angles, dtype=np.float64)
    A, Tn, Cn = angles.size, tx.size, steps.size
    stats = dict(kernel_ms=0.0, ray_steps=0, launches=0)

    def add(s):
        for k in stats:
            stats[k] += s[k]

    if group is None:
        co = eng.shoot_coarse(option, angles, src, steps, msz, (tx, ty), gamma, box, exact=exact)
        add(co.stats)
        lo, hi = brackets_from_index(co.index, angles)
        rf = eng.shoot_refine(option, lo, hi, src, steps, msz, (tx, ty), gamma, box, exact=exact)
        add(rf.stats)
        res, theta, index, cmiss, iters = rf.result, rf.theta, co.index, np.sqrt(co.miss2), rf.iterations
    else:
        import torch
        import torch.distributed as dist
        grp = None if group is True else group
        world, rank = dist.get_world_size(grp), dist.get_rank(grp)
        dev = getattr(eng, "torch_device", None) or torch.device("cuda", eng.device)
        to = lambda a: torch.as_tensor(a).to(dev)
        d_steps, d_msz, d_tx, d_ty = to(steps), to(msz), to(tx), to(ty)
        # Coarse fan sharded by angle, refinement by target -- both dealt block-cyclically in blocks of 64 (two warps of
        # neighbours).  Rays near the critical angle run to max_size, so contiguous shards hand one rank most of the
        # work (measured at N=8: 48 ms of kernels on the slowest rank against 22 ms on average); dealing single angles
        # round-robin balances the ranks but puts rays 8 launch-angle steps apart into neighbouring lanes, which costs
        # more in field-cell locality than the balance gains (55 ms).  Blocks keep a warp's rays adjacent.
        from .dist import cyclic_indices
        BLK = 64
        aidx = cyclic_indices(A, world, rank, BLK)
        if aidx.size:
            co = eng.shoot_coarse(option, to(np.ascontiguousarray(angles[aidx])), src, d_steps, d_msz, (d_tx, d_ty), gamma, box,
                                  exact=exact)
            add(co.stats)
            miss2, index = co.miss2, to(aidx)[co.index]     # local position -> index into `angles`
        else:
            miss2 = torch.full((Cn, Tn), float("inf"), dtype=torch.float64, device=dev)
            index = torch.full((Cn, Tn), INT64_MAX, dtype=torch.int64, device=dev)
        miss2, index = reduce_argmin(miss2, index, grp)   # the path's real exchange step (NCCL MIN)
        owners = [cyclic_indices(Tn, world, r, BLK) for r in range(world)]
        mine = to(owners[rank])
        lo, hi = brackets_from_index(index[:, mine].contiguous(), to(angles))
        Tl = max(o.size for o in owners)                   # targets per rank, padded to the same length for the gather
        part = torch.zeros(6, Cn, Tl, dtype=torch.float64, device=dev)
        iters = 0
        if mine.numel():
            rf = eng.shoot_refine(option, lo, hi, src, d_steps, d_msz, (d_tx[mine].contiguous(), d_ty[mine].contiguous()),
                                  gamma, box, exact=exact)
            add(rf.stats)
            part[0, :, :mine.numel()], part[1:, :, :mine.numel()] = rf.theta, rf.result
            iters = rf.iterations
        bufs = [torch.empty_like(part) for _ in range(world)]
        dist.all_gather(bufs, part, group=grp)             # result gather
        full = torch.empty(6, Cn, Tn, dtype=torch.float64, device=dev)
        for r in range(world):
            full[:, :, to(owners[r])] = bufs[r][:, :, :owners[r].size]
        it = torch.tensor([iters], dtype=torch.int64, device=dev)
        dist.all_reduce(it, op=dist.ReduceOp.MAX, group=grp)
        iters = int(it.item())
        if not device_arrays:
            full, index, miss2 = full.cpu().numpy(), index.cpu().numpy(), miss2.cpu().numpy()
            cmiss = np.sqrt(miss2)
        else:
            cmiss = miss2.sqrt()
        theta, res = full[0], full[1:]
    nsteps = res[4].astype(np.int64) if not _is_torch(res) else res[4].to(dtype=__import__("torch").int64)
    return dict(theta=theta, miss=res[0], T=res[1], exit_x=res[2], exit_y=res[3], nsteps=nsteps, index=index,
                coarse_miss=cmiss, iterations=iters, stats=stats, steps=steps, max_sizes=msz)


def sweep_two_point(result, tol_T=1e-3, hit_tol=None):
    """DELTA_S selection for the two-point problem, in the spirit of find_index (RT_bench.py:1323-1332):
    candidates are ordered fine -> coarse (as the reference's sweeps are); the deviation of a candidate is
    the largest |T - T_finest| over the targets it hits (miss <= hit_tol, default = its own step); returns
    (index, deviation[C]): the coarsest candidate such that it and every finer one deviate less than tol_T,
    or None when none / all do (same convention as find_index)."""
    T, miss, steps = np.asarray(result["T"]), np.asarray(result["miss"]), np.asarray(result["steps"])
    hit_tol = steps if hit_tol is None else np.full(steps.shape, hit_tol)
    hits = (miss <= hit_tol[:, None]) & (miss[0:1] <= hit_tol[0])
    dev = np.array([np.max(np.abs(T[c] - T[0])[hits[c]], initial=0.0) if hits[c].any() else np.inf
                    for c in range(T.shape[0])])
    ok = dev < tol_T
    if ok.all() or not ok.any():
        return None, dev
    index = None
    for i in range(len(dev)):
        if ok[:i + 1].all():
            index = i
    return index, dev
