"""Host-side mirror of RT_bench.py's hot-path interface on top of librtb.so.

`Engine` owns one C-ABI context (one GPU).  The module-level `trazar`, `search_delta`,
`n_gradient` keep the reference's names, argument meaning and return values
(RT_bench.py:766-948, 950-958, 141-156) and run on the GPU; nothing here computes the path on
the CPU.  PyTorch is used only to hand device memory / streams to the library.
"""
import ctypes as C
import math

import numpy as np

from . import _cabi
from . import constants as K
from .field import build_field, spline_arrays


def method_id(selected_func):
    """Map op1..op11 given as int, "op6"/"6", or the reference's function object (by __name__,
    trazar compares `selected_func == op7`, RT_bench.py:833) to 1..11."""
    if callable(selected_func):
        selected_func = getattr(selected_func, "__name__", "")
    s = str(selected_func).strip().lower()
    if s.startswith("op"):
        s = s[2:]
    try:
        m = int(s)
    except ValueError:
        raise ValueError("unknown step method %r (expected op1..op11)" % (selected_func,))
    if not 1 <= m <= 11:
        raise ValueError("step method must be op1..op11, got %r" % (selected_func,))
    return m


def max_size_for(user_choice, step, divisor):
    """RT_bench.py:796-799."""
    uc = K.scenario_id(user_choice)
    if uc == "2":
        return int(K.N * divisor)
    return int(np.ceil(K.constants(uc)[4] / step) + 1)


def default_fan(user_choice):
    """(theta0, x0, y0) of the scenario's default fan, RT_bench.py:809-813."""
    uc = K.scenario_id(user_choice)
    _g, R, theta_v, pos_x = K.constants(uc)[:4]
    theta0 = np.asarray(theta_v[:R], dtype=np.float64)
    if uc == "2":
        return theta0, np.full(R, float(pos_x[0])), np.full(R, float(pos_x[1]))
    return theta0, np.asarray(pos_x, dtype=np.float64), np.full(R, -2.0)


class TraceResult(dict):
    """dict with attribute access: s_ray, d_ray, errors, last_row, px_stats, stats."""
    __getattr__ = dict.__getitem__


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class DeviceSpline:
    """One of the three splines of a field fitted on the GPU (Engine.fit_field): stands where the reference passes a
    scipy RectBivariateSpline (z, grd[0], grd[1] of interpolacion(), RT_bench.py:455-458).  `.tck` / `.degrees` read
    the coefficients back in scipy's layout; the samples are kept so another engine (another GPU) can refit."""

    def __init__(self, fit, which):
        self.fit, self.which = fit, which
        self.degrees = (1, 1) if which == 0 else (3, 3)

    @property
    def tck(self):
        if self.which not in self.fit["tck"]:
            eng = self.fit["engine"]
            eng._ensure_fit(self.fit)
            self.fit["tck"][self.which] = eng.get_field(self.which)
        return self.fit["tck"][self.which]


class Engine:
    """One librtb context on one GPU."""

    def __init__(self, device=0):
        self._lib = _cabi.load()
        self._ctx = C.c_void_p()
        rc = self._lib.rtb_create(int(device), C.byref(self._ctx))
        if rc != 0:
            raise _cabi.RtbError(rc, "rtb_create(device=%d) failed: no usable CUDA device "
                                     "(raytracing_b200 has no CPU fallback)" % device)
        self.device = int(device)
        self._field_key = None
        self._keep = None

    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx:
            self._lib.rtb_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise _cabi.RtbError(rc, self._lib.rtb_last_error(self._ctx).decode())

    def use_torch_stream(self):
        """Launch on torch's current stream: torch.cuda.Event timing brackets the kernels, and the library's work is
        ordered after whatever torch has queued on the caller's tensors (a torch.zeros(...) passed as `out`, say).
        Every call that takes torch tensors does this itself."""
        import torch
        h = torch.cuda.current_stream(self.device).cuda_stream
        # torch's default stream has handle 0, which rtb_set_stream reads as "own stream": pass the
        # explicit legacy-default-stream handle (cudaStreamLegacy == 0x1) instead
        h = h if h else 1
        if getattr(self, "_torch_stream", None) != h:
            self._check(self._lib.rtb_set_stream(self._ctx, C.c_void_p(h)))
            self._torch_stream = h

    def set_field(self, z, grd):
        """Upload z (bilinear n) and grd = (dn/dy spline, dn/dx spline) as interpolacion() returns
        them (RT_bench.py:455-458); scipy objects or (tu, tv, c, ku, kv) tuples."""
        if isinstance(z, DeviceSpline):
            if not (isinstance(grd[0], DeviceSpline) and grd[0].fit is z.fit and grd[1].fit is z.fit):
                raise ValueError("set_field: z and grd must come from one fit_field call")
            return self._ensure_fit(z.fit)
        key = (id(z), id(grd[0]), id(grd[1]))
        if key == self._field_key:
            return
        arrs = [spline_arrays(s) for s in (z, grd[0], grd[1])]
        structs = [_cabi.RtbSpline(a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data,
                                   a[0].size, a[1].size, a[3], a[4]) for a in arrs]
        self._check(self._lib.rtb_set_field(self._ctx, *[C.byref(s) for s in structs]))
        self._field_key = key
        self._keep = (z, grd)  # keep ids alive so the cache key stays valid

    def _ensure_fit(self, fit):
        """Make the field of a fit_field result the resident one (refit from the kept samples if it was replaced)."""
        key = ("fit", id(fit))
        if key == self._field_key:
            return
        x, y, Z = fit["x"], fit["y"], fit["Z"]
        stats = _cabi.RtbStats()
        if _is_torch(Z):
            self.use_torch_stream()
            kind, ptrs = _cabi.RTB_PTR_DEVICE, (y.data_ptr(), x.data_ptr(), Z.data_ptr())
        else:
            kind, ptrs = _cabi.RTB_PTR_HOST, (y.ctypes.data, x.ctypes.data, Z.ctypes.data)
        self._check(self._lib.rtb_fit_field(self._ctx, kind, ptrs[0], len(y), ptrs[1], len(x), ptrs[2],
                                            float(fit["delta"]), C.byref(stats)))
        fit["stats"] = self._stats(stats)
        self._field_key = key
        self._keep = fit

    def fit_field(self, x, y, Z, delta):
        """interpolacion(x, y, Z) (RT_bench.py:435-464) on the GPU: np.gradient + the three spline fits + upload in
        one call (rtb_fit_field).  x, y, Z: numpy arrays, or torch CUDA tensors on this engine's device.
        Returns (z, (grd0, grd1)) DeviceSpline handles that trazar / search_delta / n_gradient accept."""
        if _is_torch(Z):
            import torch
            x, y, Z = (t.to(dtype=torch.float64).contiguous() for t in (x, y, Z))
            shape = tuple(Z.shape)
        else:
            x = np.ascontiguousarray(x, dtype=np.float64)
            y = np.ascontiguousarray(y, dtype=np.float64)
            Z = np.ascontiguousarray(Z, dtype=np.float64)
            shape = Z.shape
        if len(shape) != 2 or shape != (len(y), len(x)):
            raise ValueError("fit_field: Z must have shape (len(y), len(x))")    # scipy raises the same way
        fit = dict(engine=self, x=x, y=y, Z=Z, delta=float(delta), tck={})
        self._ensure_fit(fit)
        return DeviceSpline(fit, 0), (DeviceSpline(fit, 1), DeviceSpline(fit, 2))

    def gradient(self, Z, delta):
        """np.gradient(Z, delta, edge_order=2) (RT_bench.py:449) -> [d/d axis0, d/d axis1], bit-exact."""
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        g0, g1 = np.empty_like(Z), np.empty_like(Z)
        self._check(self._lib.rtb_gradient(self._ctx, _cabi.RTB_PTR_HOST, Z.ctypes.data, Z.shape[0], Z.shape[1],
                                           float(delta), g0.ctypes.data, g1.ctypes.data))
        return [g0, g1]

    def get_field(self, which):
        """(tu, tv, c) of the resident field: 0 = z, 1 = grd[0], 2 = grd[1] (scipy's .tck layout)."""
        nu, nv = C.c_int32(), C.c_int32()
        self._check(self._lib.rtb_get_field(self._ctx, which, C.byref(nu), C.byref(nv), None, None, None))
        k = 1 if which == 0 else 3
        tu, tv = np.empty(nu.value), np.empty(nv.value)
        c = np.empty((nu.value - k - 1) * (nv.value - k - 1))
        self._check(self._lib.rtb_get_field(self._ctx, which, None, None, tu.ctypes.data, tv.ctypes.data,
                                            c.ctypes.data))
        return tu, tv, c

    # -- fused multi-GPU result gather (include/rtb.h rtb_gather_*, rtb_set_gather) --
    def gather_alloc(self, nbytes):
        """(device pointer, 64-byte IPC handle) of a new buffer on this GPU that peers can map."""
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        self._check(self._lib.rtb_gather_alloc(self._ctx, int(nbytes), C.byref(ptr), handle))
        return ptr.value, handle.raw

    def gather_open(self, handle):
        ptr = C.c_void_p()
        self._check(self._lib.rtb_gather_open(self._ctx, C.create_string_buffer(bytes(handle), 64), C.byref(ptr)))
        return ptr.value

    def gather_close(self, ptr, own):
        self._check(self._lib.rtb_gather_close(self._ctx, C.c_void_p(ptr), 1 if own else 0))

    def set_gather(self, bases, R_total, offset):
        """While set, device-pointer traces also store d_ray into every buffer of `bases` (rank order, own
        included) at ray offset `offset` of a (3, R_total) array.  set_gather(None, 0, 0) switches it off."""
        n = len(bases) if bases else 0
        arr = (C.c_void_p * max(n, 1))(*[C.c_void_p(b) for b in (bases or [])])
        self._check(self._lib.rtb_set_gather(self._ctx, n, arr, int(R_total), int(offset)))

    def set_gather_cyclic(self, bases, R_total, block, nranks, rank):
        """set_gather for block-cyclic shards (dist.cyclic_indices): local ray r is global ray
        ((r // block) * nranks + rank) * block + r % block."""
        arr = (C.c_void_p * len(bases))(*[C.c_void_p(b) for b in bases])
        self._check(self._lib.rtb_set_gather_cyclic(self._ctx, len(bases), arr, int(R_total), int(block), int(nranks),
                                                    int(rank)))

    def set_option(self, name, value):
        """Tuning / test switches (include/rtb.h rtb_set_option): "queue" 0 auto | 1 always | 2 never (static grid);
        "traj_store" 0 per-lane stores (default) | 1 rows staged in shared memory + TMA stores; "shoot_groups" 0 auto | 1..16; "shoot_speculate" 0 | 1 off (default) | 2 on."""
        opt = {"queue": _cabi.RTB_OPT_QUEUE, "traj_store": _cabi.RTB_OPT_TRAJ_STORE,
               "shoot_groups": _cabi.RTB_OPT_SHOOT_GROUPS, "shoot_speculate": _cabi.RTB_OPT_SHOOT_SPECULATE}[name]
        self._check(self._lib.rtb_set_option(self._ctx, opt, int(value)))

    def build_id(self):
        return self._lib.rtb_build_id().decode()

    def kernel_id(self):
        return self._lib.rtb_kernel_id().decode()

    def sweep(self, method, theta0, x0, y0, steps, max_sizes, gamma, box, metric, skip_lo=0, skip_hi=0, exact=False):
        """rtb_sweep: the DELTA_S sweep (RT_bench.py:1303-1318) as batched traces of candidates x rays with the
        per-candidate reductions on the device.  metric: "snell" | "closure" | "px_cv".  numpy in, numpy out:
        (sum[C], max[C], count[C], stats)."""
        theta0 = np.ascontiguousarray(theta0, dtype=np.float64).ravel()
        R = theta0.size
        x0 = np.ascontiguousarray(np.broadcast_to(np.asarray(x0, dtype=np.float64), (R,)))
        y0 = np.ascontiguousarray(np.broadcast_to(np.asarray(y0, dtype=np.float64), (R,)))
        steps = np.ascontiguousarray(steps, dtype=np.float64)
        max_sizes = np.ascontiguousarray(max_sizes, dtype=np.int64)
        Cn = steps.size
        assert max_sizes.size == Cn
        sm, mx, cnt = np.zeros(Cn), np.zeros(Cn), np.zeros(Cn, dtype=np.int64)
        mid = {"snell": _cabi.RTB_SWEEP_SNELL, "closure": _cabi.RTB_SWEEP_CLOSURE, "px_cv": _cabi.RTB_SWEEP_PX_CV}[metric]
        flags = _cabi.RTB_F_EXACT_BASIS if exact else 0
        p = _cabi.RtbParams(method_id(method), flags, _cabi.RTB_PTR_HOST, 1, 1, float(gamma), 0.0,
                            float(box[0]), float(box[1]), float(box[2]), float(box[3]), None, None)
        spec = _cabi.RtbSweepSpec(steps.ctypes.data, max_sizes.ctypes.data, Cn, mid, int(skip_lo), int(skip_hi))
        stats = _cabi.RtbStats()
        self._check(self._lib.rtb_sweep(self._ctx, C.byref(p), theta0.ctypes.data, x0.ctypes.data, y0.ctypes.data, R,
                                        C.byref(spec), sm.ctypes.data, mx.ctypes.data, cnt.ctypes.data, C.byref(stats)))
        return sm, mx, cnt, self._stats(stats)

    def sweep_noalloc(self, method, theta0, x0, y0, steps, max_sizes, gamma, box, metric, sum, max, count,
                      skip_lo=0, skip_hi=0, exact=False, ptr_kind=_cabi.RTB_PTR_DEVICE):
        """rtb_sweep_noalloc: rtb_sweep on the library's fixed scratch array, for callers that cannot allocate (no
        counterpart in the reference).  Every array is a contiguous torch CUDA tensor the CALLER made -- inputs and the
        outputs sum / max (float64, C) and count (int64, C); nothing is allocated here or in the library.
        metric: "closure" | "px_cv"; C * R <= RTB_NOALLOC_MAX_RAYS.  Returns stats."""
        self.use_torch_stream()
        R, Cn = theta0.numel(), steps.numel()
        for t in (theta0, x0, y0, steps, max_sizes, sum, max, count):
            assert t.is_cuda and t.is_contiguous()
        assert x0.numel() == R and y0.numel() == R and max_sizes.numel() == Cn
        assert sum.numel() == Cn and max.numel() == Cn and count.numel() == Cn
        mid = {"snell": _cabi.RTB_SWEEP_SNELL, "closure": _cabi.RTB_SWEEP_CLOSURE, "px_cv": _cabi.RTB_SWEEP_PX_CV}[metric]
        flags = _cabi.RTB_F_EXACT_BASIS if exact else 0
        p = _cabi.RtbParams(method_id(method), flags, ptr_kind, 1, 1, float(gamma), 0.0,
                            float(box[0]), float(box[1]), float(box[2]), float(box[3]), None, None)
        spec = _cabi.RtbSweepSpec(steps.data_ptr(), max_sizes.data_ptr(), Cn, mid, int(skip_lo), int(skip_hi))
        stats = _cabi.RtbStats()
        self._check(self._lib.rtb_sweep_noalloc(self._ctx, C.byref(p), theta0.data_ptr(), x0.data_ptr(), y0.data_ptr(), R,
                                                C.byref(spec), sum.data_ptr(), max.data_ptr(), count.data_ptr(),
                                                C.byref(stats)))
        return self._stats(stats)

    def fp64_peak(self):
        v = C.c_double()
        self._check(self._lib.rtb_fp64_peak(self._ctx, C.byref(v)))
        return v.value

    def n_gradient(self, x, y, exact=False):
        """Batched n_gradient (RT_bench.py:141-156): returns n, gx, gy arrays."""
        x = np.ascontiguousarray(x, dtype=np.float64).ravel()
        y = np.ascontiguousarray(y, dtype=np.float64).ravel()
        n, gx, gy = (np.empty_like(x) for _ in range(3))
        flags = _cabi.RTB_F_EXACT_BASIS if exact else 0
        self._check(self._lib.rtb_field_eval(self._ctx, _cabi.RTB_PTR_HOST, flags, x.ctypes.data, y.ctypes.data,
                                             x.size, n.ctypes.data, gx.ctypes.data, gy.ctypes.data))
        return n, gx, gy

    def trace(self, method, theta0, x0, y0, step, max_size, gamma, box, interface=False, full=True,
              out_stride=1, want_last_row=True, want_px_stats=False, want_errors=None, exact=False,
              steps=None, max_sizes=None, out=None, tail_zero=True, static_grid=False, ray_queue=False, sort_fan=False):
        """trazar's integration for an explicit fan.

        Inputs are numpy arrays (host pointers, staged inside the call) or torch CUDA tensors
        (device pointers; outputs are then torch tensors too).  `out` may supply preallocated
        outputs by name.  Returns TraceResult.

        sort_fan (no trajectory; RTB_F_SORT_FAN): the library traces the rays in order of (x0, y0, theta0) and hands the
        results back in the caller's order.  Rays are independent, so the results are the same bits; what changes is
        that neighbouring lanes then walk through the same field cells -- an unsorted fan is 5x slower than a sorted one
        because every warp fetches 32 different cell records per step.  The ordering is three
        radix sorts on the device; it costs about as much as 10 ray-steps per ray.
        """
        m = method_id(method)
        dev = _is_torch(theta0)
        out = dict(out or {})
        if want_errors is None:
            want_errors = bool(interface)
        if dev:
            import torch
            self.use_torch_stream()
            R = theta0.numel()
            f64 = dict(dtype=torch.float64, device=theta0.device)
            def prep(a):
                if not _is_torch(a):
                    a = torch.as_tensor(np.broadcast_to(np.asarray(a, dtype=np.float64), (R,)).copy()).to(theta0.device)
                assert a.dtype == torch.float64 and a.is_contiguous() and a.numel() == R
                return a
            empty = lambda *shape: torch.empty(*shape, **f64)
            ptr = lambda a: C.c_void_p(a.data_ptr()) if a is not None else None
        else:
            theta0 = np.ascontiguousarray(theta0, dtype=np.float64).ravel()
            R = theta0.size
            prep = lambda a: np.ascontiguousarray(np.broadcast_to(np.asarray(a, dtype=np.float64), (R,)))
            empty = lambda *shape: np.empty(shape, dtype=np.float64)
            ptr = lambda a: C.c_void_p(a.ctypes.data) if a is not None else None
        theta0, x0, y0 = prep(theta0), prep(x0), prep(y0)
        max_size = int(max_size)
        rows = (max_size + out_stride - 1) // out_stride
        s_ray = out.get("s_ray", empty(rows, 6, R) if full else None)
        d_ray = out.get("d_ray", empty(3, R))
        errors = out.get("errors", empty(R) if want_errors else None)
        last_row = out.get("last_row", empty(6, R) if want_last_row else None)
        px_stats = out.get("px_stats", empty(3, R) if want_px_stats else None)
        flags = (_cabi.RTB_F_INTERFACE if interface else 0) | (_cabi.RTB_F_EXACT_BASIS if exact else 0) | \
                (0 if tail_zero else _cabi.RTB_F_NO_TAIL_ZERO) | (_cabi.RTB_F_STATIC_GRID if static_grid else 0) | (_cabi.RTB_F_RAY_QUEUE if ray_queue else 0) | \
                (_cabi.RTB_F_SORT_FAN if sort_fan else 0)
        steps_a = max_sizes_a = None
        if steps is not None:
            if dev:
                steps_a = steps
                max_sizes_a = max_sizes
                sp, mp = ptr(steps_a), ptr(max_sizes_a)
            else:
                steps_a = np.ascontiguousarray(steps, dtype=np.float64)
                max_sizes_a = np.ascontiguousarray(max_sizes, dtype=np.int64)
                sp, mp = ptr(steps_a), ptr(max_sizes_a)
        else:
            sp = mp = None
        p = _cabi.RtbParams(m, flags, _cabi.RTB_PTR_DEVICE if dev else _cabi.RTB_PTR_HOST, int(out_stride),
                            max_size, float(gamma), float(step if step is not None else 0.0),
                            float(box[0]), float(box[1]), float(box[2]), float(box[3]), sp, mp)
        stats = _cabi.RtbStats()
        self._check(self._lib.rtb_trace(self._ctx, C.byref(p), ptr(theta0), ptr(x0), ptr(y0), R, ptr(s_ray),
                                        ptr(d_ray), ptr(errors), ptr(last_row), ptr(px_stats), C.byref(stats)))
        return TraceResult(s_ray=s_ray, d_ray=d_ray, errors=errors, last_row=last_row, px_stats=px_stats,
                           stats=self._stats(stats))

    # -- batched two-point shooting search (include/rtb.h rtb_shoot_*; parity-unpinned) --
    def _shoot_args(self, method, src, steps, max_sizes, targets, gamma, box, exact, like):
        """Common marshalling: arrays follow `like` (torch CUDA tensor -> device pointers, else host)."""
        dev = _is_torch(like)
        if dev:
            import torch
            self.use_torch_stream()
            f64 = lambda a: a.contiguous() if _is_torch(a) else torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).to(like.device)
            i64 = lambda a: a if _is_torch(a) else torch.as_tensor(np.ascontiguousarray(a, dtype=np.int64)).to(like.device)
            ptr = lambda a: C.c_void_p(a.data_ptr()) if a is not None else None
            empty = lambda shape, dt=None: torch.empty(*shape, dtype=dt or torch.float64, device=like.device)
            i64t = torch.int64
        else:
            f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
            i64 = lambda a: np.ascontiguousarray(a, dtype=np.int64)
            ptr = lambda a: C.c_void_p(a.ctypes.data) if a is not None else None
            empty = lambda shape, dt=None: np.empty(shape, dtype=dt or np.float64)
            i64t = np.int64
        steps, max_sizes = f64(steps), i64(max_sizes)
        tx, ty = f64(targets[0]), f64(targets[1])
        Cn, Tn = int(steps.shape[0]), int(tx.shape[0])
        flags = _cabi.RTB_F_EXACT_BASIS if exact else 0
        p = _cabi.RtbParams(method_id(method), flags, _cabi.RTB_PTR_DEVICE if dev else _cabi.RTB_PTR_HOST, 1, 1,
                            float(gamma), 0.0, float(box[0]), float(box[1]), float(box[2]), float(box[3]), None, None)
        q = _cabi.RtbShootProblem(float(src[0]), float(src[1]), ptr(steps), ptr(max_sizes), Cn, ptr(tx), ptr(ty), Tn)
        keep = (steps, max_sizes, tx, ty)
        return p, q, Cn, Tn, f64, ptr, empty, i64t, keep

    @staticmethod
    def _stats(stats):
        return dict(kernel_ms=stats.kernel_ms, total_ms=stats.total_ms, ray_steps=stats.ray_steps,
                    launches=stats.launches, h2d_bytes=stats.h2d_bytes, d2h_bytes=stats.d2h_bytes,
                    nonfinite_rays=stats.nonfinite_rays)

    def shoot_coarse(self, method, angles, src, steps, max_sizes, targets, gamma, box, index0=0, want_exit=False,
                     exact=False):
        """Coarse phase: trace the fan `angles` for every DELTA_S candidate and bracket every target.
        Returns TraceResult(miss2 (C,Tn), index (C,Tn) int64, exit (3,C,A) or None, stats)."""
        p, q, Cn, Tn, f64, ptr, empty, i64t, keep = self._shoot_args(method, src, steps, max_sizes, targets, gamma, box,
                                                                     exact, angles)
        angles = f64(angles)
        A = int(angles.shape[0])
        miss2, index = empty((Cn, Tn)), empty((Cn, Tn), i64t)
        ex = empty((3, Cn, A)) if want_exit else None
        stats = _cabi.RtbStats()
        self._check(self._lib.rtb_shoot_coarse(self._ctx, C.byref(p), C.byref(q), ptr(angles), A, int(index0), ptr(ex),
                                               ptr(miss2), ptr(index), C.byref(stats)))
        return TraceResult(miss2=miss2, index=index, exit=ex, stats=self._stats(stats))

    def shoot_refine(self, method, lo, hi, src, steps, max_sizes, targets, gamma, box, exact=False):
        """golden(miss^2, lo, hi) (RT_bench.py:175-199) for every (candidate, target) pair.
        Returns TraceResult(theta (C,Tn), result (5,C,Tn) = miss, T, exit x, exit y, steps; iterations; stats)."""
        p, q, Cn, Tn, f64, ptr, empty, i64t, keep = self._shoot_args(method, src, steps, max_sizes, targets, gamma, box,
                                                                     exact, lo)
        lo, hi = f64(lo), f64(hi)
        assert tuple(lo.shape) == (Cn, Tn) and tuple(hi.shape) == (Cn, Tn)
        theta, result = empty((Cn, Tn)), empty((5, Cn, Tn))
        it = C.c_int64(0)
        stats = _cabi.RtbStats()
        self._check(self._lib.rtb_shoot_refine(self._ctx, C.byref(p), C.byref(q), ptr(lo), ptr(hi), ptr(theta),
                                               ptr(result), C.byref(it), C.byref(stats)))
        return TraceResult(theta=theta, result=result, iterations=it.value, stats=self._stats(stats))
