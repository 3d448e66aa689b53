"""ctypes loader for the CPU oracle (oracle/lsmc_oracle.cpp).  TEST INFRASTRUCTURE: importable only from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs — never from storage_b200/."""
import ctypes as C
import os
import subprocess

import numpy as np

from storage_b200 import _abi  # struct definitions of include/lsmc_b200.h only; no product code is executed

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liblsmc_oracle.so")


class oracle_options(C.Structure):
    _fields_ = [("regression_mode", C.c_int32), ("num_threads", C.c_int32)]


class oracle_panels(C.Structure):
    _fields_ = [(n, _abi.c_double_p) for n in ("inventory", "inject_withdraw_volume", "cmdty_consumed", "inventory_loss",
                                               "net_volume", "pv", "decision_index", "backward_value_first")]


def build():
    subprocess.run(["make", "-s", "-C", _HERE], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_last_error.restype = C.c_char_p
    return _lib


class OracleError(Exception):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def _check(code):
    if code != 0:
        raise OracleError(code, lib().oracle_last_error().decode())


def _dp(a):
    return a.ctypes.data_as(_abi.c_double_p)


def grid_points(lower, upper, spacing):
    cnt = C.c_int()
    out = np.zeros(100000)
    _check(lib().oracle_grid_points(C.c_double(lower), C.c_double(upper), C.c_double(spacing), _dp(out), len(out), C.byref(cnt)))
    return out[:cnt.value].copy()


def bang_bang_decision_set(min_rate, max_rate, inventory, inventory_loss, next_min, next_max, tol, extra=0):
    out = np.zeros(64)
    cnt = C.c_int()
    _check(lib().oracle_bang_bang_decision_set(C.c_double(min_rate), C.c_double(max_rate), C.c_double(inventory),
                                               C.c_double(inventory_loss), C.c_double(next_min), C.c_double(next_max),
                                               C.c_double(tol), C.c_int(extra), _dp(out), C.byref(cnt)))
    return out[:cnt.value].copy()


def bisect_inventory_space(grid, inventory, tol):
    g = np.ascontiguousarray(grid, dtype=np.float64)
    lo, hi = C.c_int(), C.c_int()
    _check(lib().oracle_bisect_inventory_space(_dp(g), len(g), C.c_double(inventory), C.c_double(tol), C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


def max_value_and_index(values):
    a = np.ascontiguousarray(values, dtype=np.float64)
    mx, idx = C.c_double(), C.c_int()
    _check(lib().oracle_max_value_and_index(_dp(a), len(a), C.byref(mx), C.byref(idx)))
    return mx.value, idx.value


def calculate_inventory_space(storage_struct, starting_inventory):
    P = storage_struct.num_periods
    mn, mx = np.zeros(P), np.zeros(P)
    _check(lib().oracle_calculate_inventory_space(C.byref(storage_struct), C.c_double(starting_inventory), _dp(mn), _dp(mx)))
    return mn, mx


def inject_withdraw_range(storage_struct, period, inventory):
    mn, mx = C.c_double(), C.c_double()
    _check(lib().oracle_inject_withdraw_range(C.byref(storage_struct), period, C.c_double(inventory), C.byref(mn), C.byref(mx)))
    return mn.value, mx.value


def constraint_space_bounds(storage_struct, period, next_lo, next_hi):
    lo, hi = C.c_double(), C.c_double()
    _check(lib().oracle_constraint_space_bounds(C.byref(storage_struct), period, C.c_double(next_lo), C.c_double(next_hi),
                                                C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


def populate_design_matrix(powers, spot, factors):
    powers = np.ascontiguousarray(powers, dtype=np.int32)
    K, F = powers.shape[0], powers.shape[1] - 1
    spot = np.ascontiguousarray(spot, dtype=np.float64)
    facs = [np.ascontiguousarray(f, dtype=np.float64) for f in factors]
    N = spot.shape[0]
    x = np.zeros((K, N))
    ptrs = (C.c_void_p * max(F, 1))(*[f.ctypes.data for f in facs])
    _check(lib().oracle_populate_design_matrix(powers.ctypes.data_as(_abi.c_int32_p), K, F, _dp(spot), ptrs, C.c_int64(N), _dp(x)))
    return x  # [K][N] = column-major N x K


def regress(x_kn, y, mode):
    x = np.ascontiguousarray(x_kn, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    K, N = x.shape
    fitted, coeffs = np.zeros(N), np.zeros(K)
    _check(lib().oracle_regress(_dp(x), K, C.c_int64(N), _dp(y), mode, _dp(fitted), _dp(coeffs)))
    return fitted, coeffs


class OracleResults:
    pass


def lsmc_calculate(flat, reg_spot, reg_factors, val_spot, val_factors, mode=1, threads=0, panels=True):
    """oracle_lsmc_calculate on a storage_b200.FlatProblem and [rows][N] simulation arrays."""
    pb = flat.struct
    P, N = flat.num_periods, int(pb.num_sims)
    reg_spot = np.ascontiguousarray(reg_spot, dtype=np.float64)
    val_spot = np.ascontiguousarray(val_spot, dtype=np.float64)
    regf = [np.ascontiguousarray(f, dtype=np.float64) for f in reg_factors]
    valf = [np.ascontiguousarray(f, dtype=np.float64) for f in val_factors]
    rp = (C.c_void_p * max(len(regf), 1))(*[f.ctypes.data for f in regf])
    vp = (C.c_void_p * max(len(valf), 1))(*[f.ctypes.data for f in valf])
    out = OracleResults()
    out.deltas = np.zeros(P + 1)
    out.profile = np.zeros((P + 1, 5))
    out.trigger = (_abi.lsmc_trigger_prices * max(P, 1))()
    out.pv_by_sim = np.zeros(N)
    r = _abi.lsmc_results()
    r.deltas, r.profile, r.trigger_prices, r.pv_by_sim = _dp(out.deltas), _dp(out.profile), out.trigger, _dp(out.pv_by_sim)
    pn = oracle_panels()
    if panels:
        for name in ("inventory", "inject_withdraw_volume", "cmdty_consumed", "inventory_loss", "net_volume", "pv"):
            arr = np.zeros((P + 1, N))
            setattr(out, name, arr)
            setattr(pn, name, _dp(arr))
        out.decision_index = np.zeros((P, N))
        pn.decision_index = _dp(out.decision_index)
        out.backward_value_first = np.zeros(N)
        pn.backward_value_first = _dp(out.backward_value_first)
    opt = oracle_options(mode, threads)
    _check(lib().oracle_lsmc_calculate(C.byref(pb), _dp(reg_spot), rp, _dp(val_spot), vp, C.byref(opt), C.byref(r), C.byref(pn)))
    out.npv, out.std_error, out.backward_npv = r.npv, r.std_error, r.backward_npv
    out.snapped_inventories = int(r.snapped_inventories)  # forward-simulation inventories moved onto a limit (documented departure)
    return out


def intrinsic_calculate(flat):
    """oracle_intrinsic_calculate on a storage_b200.FlatProblem: (npv, profile[P+1][5])."""
    P = flat.num_periods
    npv = C.c_double()
    profile = np.zeros((P + 1, 5))
    _check(lib().oracle_intrinsic_calculate(C.byref(flat.struct), C.byref(npv), _dp(profile)))
    return npv.value, profile


def num_threads():
    return lib().oracle_num_threads()
