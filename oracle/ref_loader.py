"""TEST INFRASTRUCTURE ONLY -- loader for the *unmodified* reference (RT_bench.py).

Only tests/, tests/golden/gen_golden.py and ad-hoc validation may use this.  It works
only where /root/reference exists (the build container); the GPU box has no copy, so
nothing in `-m gpu` tests, smoke() or bench.py imports this module.

RT_bench.py imports matplotlib at module top (RT_bench.py:43-50) which is not installed,
so stub modules are inserted into sys.modules first; the `__main__` guard (RT_bench.py:1550)
keeps the interactive menus from running.  Module globals are then set the way `__main__`
does (RT_bench.py:1567-1584).
"""
import importlib.util
import os
import sys
import types
import warnings

import numpy as np

REF_PATH = os.environ.get("RT_REFERENCE", "/root/reference/RT_bench.py")

_SCENARIO_F = {"1": "interface", "2": "fisheye", "3": "vert_heterogeneous", "4": "vert_heterogeneous"}


def available():
    return os.path.exists(REF_PATH)


def _stub_matplotlib():
    def mod(name, **attrs):
        m = sys.modules.get(name)
        if m is None:
            m = types.ModuleType(name)
            sys.modules[name] = m
        for k, v in attrs.items():
            setattr(m, k, v)
        return m

    if "matplotlib" in sys.modules and not getattr(sys.modules["matplotlib"], "_rt_stub", False):
        return
    mpl = mod("matplotlib", _rt_stub=True)
    mpl.pyplot = mod("matplotlib.pyplot")
    mpl.animation = mod("matplotlib.animation", FuncAnimation=object)
    mpl.ticker = mod("matplotlib.ticker", FormatStrFormatter=object)
    tk = mod("mpl_toolkits")
    tk.axes_grid1 = mod("mpl_toolkits.axes_grid1", make_axes_locatable=lambda *a, **k: None)


def load_reference():
    """Import RT_bench.py as module `RT` (fresh instance each call)."""
    _stub_matplotlib()
    spec = importlib.util.spec_from_file_location("RT_reference", REF_PATH)
    RT = importlib.util.module_from_spec(spec)
    sys.dont_write_bytecode = True  # /root/reference is read-only
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", SyntaxWarning)  # RT_bench.py:1188 LaTeX escapes
        spec.loader.exec_module(RT)
    warnings.filterwarnings("ignore", category=DeprecationWarning)  # np.cross on 2-vectors (RT_bench.py:360)
    return RT


def setup_scenario(RT, user_choice, fan=None):
    """Do what `__main__` does for scenario `user_choice` ("1".."4"), RT_bench.py:1565-1588.

    fan: optional dict(theta_v=..., pos_x=..., ray_count=...) overriding the scenario's
    default ray fan by wrapping RT.constants (trazar re-reads constants(), RT_bench.py:793).
    Returns (z, grd).
    """
    user_choice = str(user_choice)
    RT.f = getattr(RT, _SCENARIO_F[user_choice])
    if fan is not None:
        base = RT.constants

        def patched(uc, _base=base, _fan=fan):
            out = list(_base(uc))
            theta = np.asarray(_fan["theta_v"], dtype=np.float64)
            out[1] = int(_fan.get("ray_count", len(theta)))
            out[2] = theta
            if "pos_x" in _fan:
                out[3] = np.asarray(_fan["pos_x"], dtype=np.float64)
            elif uc != "2":
                out[3] = np.ones(out[1]) * -2
            return tuple(out)

        RT.constants = patched
    (RT.gamma, RT.ray_count, RT.theta_v, RT.pos_x, RT.s, RT.limx_i, RT.limx_s, RT.limy_i, RT.limy_s,
     RT.op_interface, RT.op_fish, RT.op_vert_heterogeneous, RT.op_anisotropy) = RT.constants(user_choice)
    linx, liny, X, Y, Z = RT.genZ(RT.limx_i, RT.limx_s, RT.limy_i, RT.limy_s)
    z, grd, _hess = RT.interpolacion(linx, liny, Z, X, Y)
    return z, grd


def method(RT, m):
    return getattr(RT, "op%d" % int(m))
