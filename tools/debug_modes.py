"""Dev: where do the output modes of one scenario / method differ?  usage: debug_modes.py <scenario> <method>"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
import raytracing_b200 as rt
from _util import scenario_setup
uc, m = sys.argv[1], int(sys.argv[2])
s = scenario_setup(uc)
eng = rt.get_engine(0)
eng.set_field(s["z"], s["grd"])
a = (s["theta0"], s["x0"], s["y0"], s["step"], s["M"], s["gamma"], s["box"])
full = eng.trace(m, *a, full=True)
fin = eng.trace(m, *a, full=False)
st = eng.trace(m, *a, full=True, out_stride=7)
print("d_ray equal", np.array_equal(full.d_ray, fin.d_ray), "last_row equal", np.array_equal(full.last_row, fin.last_row))
print("steps", full.d_ray[2][:8], fin.d_ray[2][:8])
d = np.abs(full.last_row - fin.last_row); print("last_row max diff", d.max(), "rays differing", np.nonzero(d.max(axis=0))[0])
print("per column (x y px py T ang):", d.max(axis=1))
d = np.abs(st.s_ray - full.s_ray[::7]); print("strided max diff", d.max())
bad = np.nonzero(d.max(axis=(0, 1)))[0]; print("rays differing (strided)", bad)
for r in bad[:3]:
    rows = np.nonzero(d[:, :, r].max(axis=1))[0]
    print(" ray", r, "first differing kept row", rows[:3], "diff", d[rows[0], :, r])
full2 = eng.trace(m, *a, full=True)
print("full vs full again equal:", np.array_equal(full.s_ray, full2.s_ray))
