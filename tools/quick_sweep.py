import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_b200 as rt
t = time.time()
z, grd = rt.build_field("2")
r = rt.find_delta_s(6, z, grd, "2")
print("fisheye sweep index", r["index"], r["divisor"], time.time() - t)
