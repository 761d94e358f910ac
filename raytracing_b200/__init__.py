"""raytracing_b200 -- B200-native (sm_100a) replacement for the ray-stepping hot path of
neyuru/RayTracing's RT_bench.py: trazar / search_delta / n_gradient, the eleven step methods,
the DELTA_S finder and the benchmark statistic, behind the C ABI of include/rtb.h.
"""
from . import constants
from .constants import (SIGMA, DELTA, DELTA_S, DELTA_S_DIVISOR_FISHEYE, N, GOLD_TOL, anisotropy,
                        constants as scenario_constants, calibrated_step, F_ALG)
from .field import genZ, interpolacion, build_field
from .engine import Engine, DeviceSpline, method_id, max_size_for, default_fan
from .fans import make_fan
from .api import (interpolacion_gpu, build_field_gpu, trazar, search_delta, n_gradient, find_delta_s, sweep_candidates, sweep_metrics,
                  find_index_interface, find_index_fisheye, find_index_vert, px_cv,
                  remove_outliers_iqr, benchmark_statistic, get_engine)
from .shooting import (shooting_search, targets_from_angles, sweep_two_point, brackets_from_index,
                       reduce_argmin, candidate_rows, default_source)
from .wavefront import wavefront_points, wavefront, pchip_knot_derivatives

__all__ = [n for n in dir() if not n.startswith("_")]
