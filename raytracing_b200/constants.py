"""Constants and scenario table of RT_bench.py, restated (values are part of the contract).

RT_bench.py:59-97 (constants), :106-119 (scenario closures), :247-295 (constants()),
:1408-1455 (calibrated DELTA_S tables).
"""
import sys

import numpy as np

# 'interface' scenario parameters RT_bench.py:59-61
THCK_PARAM = 0.005
A = (1 + np.sqrt(2)) / 2 - 99 * (np.sqrt(2) - 1) / 200
SIGMA = -2 * THCK_PARAM * np.log((A - 1) / (np.sqrt(2) - A))

# golden search RT_bench.py:64-66
DELTA_G = np.pi / 2
GOLD_RATIO = (np.sqrt(5) - 1) / 2
GOLD_TOL = np.sqrt(sys.float_info.epsilon)

MAX_DEVIATION = 0.2  # RT_bench.py:69

# grid and step RT_bench.py:77-84
DELTA = SIGMA / 3
DELTA_S_DIVISOR = 20
DELTA_S = SIGMA / DELTA_S_DIVISOR
N = 10
DELTA_S_DIVISOR_FISHEYE = 90

# DELTA_S search ranges RT_bench.py:89-97
DELTA_STEP = 0.01
DELTA_S_DIVISOR_UPPER_LIMIT = 3
DELTA_S_DIVISOR_LOWER_LIMIT = 1 + DELTA_STEP
DELTA_STEP_FISHEYE = 1
DELTA_S_DIVISOR_FISHEYE_UPPER_LIMIT = 303
DELTA_S_DIVISOR_FISHEYE_LOWER_LIMIT = 4
DELTA_STEP_VERT = 0.005
DELTA_S_DIVISOR_VERT_UPPER_LIMIT = 2
DELTA_S_DIVISOR_VERT_LOWER_LIMIT = 1 / 40


def interface(a, b):  # RT_bench.py:106-108
    with np.errstate(over="ignore"):
        return np.sqrt(2) - (np.sqrt(2) - 1) / (1 + np.exp(-b / THCK_PARAM))


def fisheye(a, b):  # RT_bench.py:110-112
    return 1 / (1 + np.power(a, 2) + np.power(b, 2))


def vert_heterogeneous(a, b):  # RT_bench.py:114-116
    return 1 / (18 + 2 * b)


def anisotropy(theta, gamma):  # RT_bench.py:118-119
    return np.sqrt((gamma * np.sin(theta)) ** 2 + np.cos(theta) ** 2)


SCENARIO_NAMES = {"1": "interface", "2": "fisheye", "3": "vert_heterogeneous", "4": "anisotropy"}
SCENARIO_IDS = {v: k for k, v in SCENARIO_NAMES.items()}
SCENARIO_FUNCS = {"1": interface, "2": fisheye, "3": vert_heterogeneous, "4": vert_heterogeneous}  # RT:1567-1579


def scenario_id(user_choice):
    """Accept "1".."4", 1..4 or the scenario name."""
    s = str(user_choice)
    s = SCENARIO_IDS.get(s, s)
    if s not in SCENARIO_NAMES:
        raise ValueError("scenario must be 1..4 or one of %s" % sorted(SCENARIO_IDS))
    return s


def constants(user_choice):
    """RT_bench.py:247-295, same 13-tuple."""
    user_choice = scenario_id(user_choice)
    if user_choice == "1":
        gamma, ray_count = 1, 42
        theta_v = np.linspace(2 * (np.pi / 60), np.pi / 2, ray_count + 1)
        pos_x = np.ones((ray_count)) * -2
        s = 80
        lim = (-2, 20, -2, 4)
        ops = (1, 0, 0, 0)
    elif user_choice == "2":
        gamma, ray_count = 1, 1
        theta_v = np.linspace(np.pi / 2, np.pi / 2, 1)
        pos_x = np.array((1, 0))
        s = N * (2 * np.pi)
        lim = (-1.5, 1.5, -1.5, 1.5)
        ops = (0, 1, 0, 0)
    else:
        gamma = 1 if user_choice == "3" else 3
        ray_count = 31
        theta_v = np.linspace(0, np.pi / 2, ray_count)
        pos_x = np.ones((ray_count)) * -2
        s = 80
        lim = (-2, 5, -2.5, 1)
        ops = (0, 0, 1, 0) if user_choice == "3" else (0, 0, 0, 1)
    return (gamma, ray_count, theta_v, pos_x, s) + lim + ops


# Calibrated DELTA_S divisors RT_bench.py:1413-1455 (method id -> divisor)
CALIBRATED_SIGMA_DIVISOR = {1: 38.64, 2: 38.37, 3: 2.34, 4: 2.53, 5: 2.53, 6: 2.55, 7: 30.05, 8: 2.74, 9: 2.74}
CALIBRATED_FISHEYE_DIVISOR = {1: 4587, 2: 4556, 3: 278, 4: 300, 5: 300, 6: 303, 7: 3567, 8: 325, 9: 325}
CALIBRATED_ANISO_DIVISOR = {10: 2.53, 11: 2.74}

# algorithmic flop per ray-step of each method (roofline numerator)
F_ALG = {1: 358, 2: 362, 3: 433, 4: 429, 5: 4846, 6: 376, 7: 376, 8: 372, 9: 4789, 10: 7614, 11: 7557}


def calibrated_step(user_choice, method):
    """DELTA_S (and fisheye divisor) the 'calibrated' menu answer selects, RT_bench.py:1411-1456.

    Returns (DELTA_S, DELTA_S_DIVISOR_FISHEYE)."""
    uc = scenario_id(user_choice)
    method = int(method)
    if uc in ("1", "3"):
        return SIGMA / CALIBRATED_SIGMA_DIVISOR[method], DELTA_S_DIVISOR_FISHEYE
    if uc == "2":
        d = CALIBRATED_FISHEYE_DIVISOR[method]
        return 2 * np.pi / d, d
    return SIGMA / CALIBRATED_ANISO_DIVISOR[method], DELTA_S_DIVISOR_FISHEYE
