"""Two implementations of the sim-independent helpers behind one interface, so the reference's unit-test vectors
(ported in test_helpers_kat.py) pin both the CPU oracle and the product library's host code."""
import ctypes as C

import numpy as np

from oracle import oracle
from storage_b200 import _abi
from storage_b200.errors import check_status


class OracleImpl:
    name = "oracle"
    ArgumentError = oracle.OracleError

    def grid_points(self, lo, hi, spacing):
        return list(oracle.grid_points(lo, hi, spacing))

    def decision_set(self, min_rate, max_rate, inv, loss, nmin, nmax, tol, extra):
        return list(oracle.bang_bang_decision_set(min_rate, max_rate, inv, loss, nmin, nmax, tol, extra))

    def bisect(self, grid, inv, tol):
        return oracle.bisect_inventory_space(grid, inv, tol)

    def inventory_space(self, storage_struct, inv):
        return oracle.calculate_inventory_space(storage_struct, inv)

    def range(self, storage_struct, period, inv):
        return oracle.inject_withdraw_range(storage_struct, period, inv)

    def space_bounds(self, storage_struct, period, next_lo, next_hi):
        return oracle.constraint_space_bounds(storage_struct, period, next_lo, next_hi)


class LibImpl:
    name = "lib"

    def __init__(self):
        self.lib = _abi.load_library()
        from storage_b200.errors import ArgumentException
        self.ArgumentError = ArgumentException

    def grid_points(self, lo, hi, spacing):
        n = self.lib.lsmc_grid_points(lo, hi, spacing, None, 0)
        if n < 0:
            check_status(-n, None)
        out = np.zeros(n)
        self.lib.lsmc_grid_points(lo, hi, spacing, out.ctypes.data_as(_abi.c_double_p), n)
        return list(out)

    def decision_set(self, min_rate, max_rate, inv, loss, nmin, nmax, tol, extra):
        out = (C.c_double * 64)()
        n = self.lib.lsmc_bang_bang_decision_set(min_rate, max_rate, inv, loss, nmin, nmax, tol, extra, out)
        if n < 0:
            check_status(-n, None)
        return list(out[:n])

    def bisect(self, grid, inv, tol):
        g = np.ascontiguousarray(grid, dtype=np.float64)
        lo, hi = C.c_int32(), C.c_int32()
        check_status(self.lib.lsmc_bisect_inventory_space(g.ctypes.data_as(_abi.c_double_p), len(g), inv, tol,
                                                          C.byref(lo), C.byref(hi)), None)
        return lo.value, hi.value

    def inventory_space(self, storage_struct, inv):
        P = storage_struct.num_periods
        mn, mx = np.zeros(P), np.zeros(P)
        check_status(self.lib.lsmc_calculate_inventory_space(C.byref(storage_struct), inv, mn.ctypes.data_as(_abi.c_double_p),
                                                             mx.ctypes.data_as(_abi.c_double_p)), None)
        return mn, mx

    def range(self, storage_struct, period, inv):
        mn, mx = C.c_double(), C.c_double()
        check_status(self.lib.lsmc_inject_withdraw_range(C.byref(storage_struct), period, inv, C.byref(mn), C.byref(mx)), None)
        return mn.value, mx.value

    def space_bounds(self, storage_struct, period, next_lo, next_hi):
        lo, hi = C.c_double(), C.c_double()
        check_status(self.lib.lsmc_inventory_space_bounds(C.byref(storage_struct), period, next_lo, next_hi,
                                                          C.byref(lo), C.byref(hi)), None)
        return lo.value, hi.value


def make(name):
    return OracleImpl() if name == "oracle" else LibImpl()
