"""ctypes view of include/lsmc_b200.h and loader of the in-tree CUDA library.

The product path has no CPU fallback: if liblsmc_b200.so is missing or cannot be loaded this module raises,
and every compute entry point fails with LSMC_ERR_CUDA when no B200 is present.
"""
import ctypes as C
import os

LSMC_MAX_BASIS = 16
LSMC_MAX_FACTORS = 8
LSMC_MAX_DECISIONS = 16
LSMC_NUM_TRIGGER_VOLUMES = 10
LSMC_COMM_ID_BYTES = 128

(LSMC_OK, LSMC_ERR_ARGUMENT, LSMC_ERR_INVENTORY_CONSTRAINTS, LSMC_ERR_INVALID_OPERATION, LSMC_ERR_CANCELLED,
 LSMC_ERR_CUDA, LSMC_ERR_NUMERICAL, LSMC_ERR_UNSUPPORTED) = range(8)

(LSMC_CONSTRAINT_CONSTANT, LSMC_CONSTRAINT_PIECEWISE_LINEAR, LSMC_CONSTRAINT_STEP,
 LSMC_CONSTRAINT_POLYNOMIAL) = range(4)

(LSMC_PANEL_INVENTORY, LSMC_PANEL_INJECT_WITHDRAW_VOLUME, LSMC_PANEL_CMDTY_CONSUMED, LSMC_PANEL_INVENTORY_LOSS,
 LSMC_PANEL_NET_VOLUME, LSMC_PANEL_PV, LSMC_PANEL_SPOT_REGRESSION, LSMC_PANEL_SPOT_VALUATION,
 LSMC_PANEL_FACTORS_REGRESSION, LSMC_PANEL_FACTORS_VALUATION, LSMC_PANEL_DECISION_INDEX) = range(11)

LSMC_RETURN_DECISION_INDEX = 1 << 16
LSMC_REFERENCE_OPERATION_ORDER = 1 << 17
LSMC_STREAM_PANELS = 1 << 18

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)


class lsmc_storage(C.Structure):
    _fields_ = [
        ("num_periods", C.c_int32),
        ("must_be_empty_at_end", C.c_int32),
        ("min_inventory", c_double_p),
        ("max_inventory", c_double_p),
        ("constraint_kind", c_int32_p),
        ("ratchet_begin", c_int32_p),
        ("ratchet_count", c_int32_p),
        ("num_ratchet_rows", C.c_int32),
        ("ratchet_inventory", c_double_p),
        ("ratchet_min_rate", c_double_p),
        ("ratchet_max_rate", c_double_p),
        ("injection_cost", c_double_p),
        ("injection_cost_df", c_double_p),
        ("withdrawal_cost", c_double_p),
        ("withdrawal_cost_df", c_double_p),
        ("consumed_on_inject", c_double_p),
        ("consumed_on_withdraw", c_double_p),
        ("inventory_loss", c_double_p),
        ("inventory_cost", c_double_p),
        ("inventory_cost_df", c_double_p),
        ("terminal_npv_coeff", C.c_double * 4),
        ("newton_raphson_accuracy", C.c_double),
        ("newton_raphson_max_iterations", C.c_int32),
        ("newton_raphson_subdivision", C.c_int32),
    ]


class lsmc_problem(C.Structure):
    _fields_ = [
        ("storage", lsmc_storage),
        ("inventory", C.c_double),
        ("current_period_is_active", C.c_int32),
        ("forward_price", c_double_p),
        ("settle_df", c_double_p),
        ("grid_spacing", C.c_double),
        ("numerical_tolerance", C.c_double),
        ("extra_decisions", C.c_int32),
        ("discount_deltas", C.c_int32),
        ("simulation_data_returned", C.c_uint32),
        ("num_factors", C.c_int32),
        ("num_basis", C.c_int32),
        ("basis_powers", c_int32_p),
        ("num_sims", C.c_int64),
        ("reference_compat_terminal_pv", C.c_int32),
    ]


class lsmc_multi_factor_params(C.Structure):
    _fields_ = [
        ("num_factors", C.c_int32),
        ("mean_reversion", c_double_p),
        ("factor_vol", c_double_p),
        ("factor_corr", c_double_p),
        ("period_time", c_double_p),
    ]


class lsmc_trigger_prices(C.Structure):
    _fields_ = [
        ("num_inject", C.c_int32),
        ("num_withdraw", C.c_int32),
        ("inject_volume", C.c_double * LSMC_NUM_TRIGGER_VOLUMES),
        ("inject_price", C.c_double * LSMC_NUM_TRIGGER_VOLUMES),
        ("withdraw_volume", C.c_double * LSMC_NUM_TRIGGER_VOLUMES),
        ("withdraw_price", C.c_double * LSMC_NUM_TRIGGER_VOLUMES),
        ("has_inject_price", C.c_int32),
        ("has_withdraw_price", C.c_int32),
        ("max_inject_volume", C.c_double),
        ("max_inject_trigger_price", C.c_double),
        ("max_withdraw_volume", C.c_double),
        ("max_withdraw_trigger_price", C.c_double),
    ]


class lsmc_results(C.Structure):
    _fields_ = [
        ("npv", C.c_double),
        ("std_error", C.c_double),
        ("backward_npv", C.c_double),
        ("deltas", c_double_p),
        ("profile", c_double_p),
        ("trigger_prices", C.POINTER(lsmc_trigger_prices)),
        ("pv_by_sim", c_double_p),
        ("ms_total", C.c_double),
        ("ms_regression_sim", C.c_double),
        ("ms_valuation_sim", C.c_double),
        ("ms_backward", C.c_double),
        ("ms_pseudo_inverse", C.c_double),
        ("ms_forward", C.c_double),
        ("kernel_launches", C.c_int64),
        ("ms_backward_kernel", C.c_double),
        ("backward_kernel_launches", C.c_int64),
        ("ms_device", C.c_double),
        ("snapped_inventories", C.c_int64),
        ("ms_upload", C.c_double),
    ]


PROGRESS_FN = C.CFUNCTYPE(None, C.c_double, C.c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LSMC_LIBRARY") or os.path.join(_HERE, "lib", "liblsmc_b200.so")  # override: tuning builds only

# every symbol include/lsmc_b200.h declares: (name, restype, argtypes)
_H = C.c_void_p
SYMBOLS = [
    ("lsmc_create", C.c_int, [C.POINTER(lsmc_problem), C.c_int, C.POINTER(_H)]),
    ("lsmc_set_sims", C.c_int, [_H, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    ("lsmc_set_sims_device", C.c_int, [_H, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    ("lsmc_set_sims_async", C.c_int, [_H, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    ("lsmc_set_panel_sink", C.c_int, [_H, C.c_int, C.c_void_p]),
    ("lsmc_simulate", C.c_int, [_H, C.c_int, C.POINTER(lsmc_multi_factor_params), C.c_uint64, C.c_int, C.c_int64]),
    ("lsmc_simulate_noalloc", C.c_int,
     [C.c_int, C.POINTER(lsmc_multi_factor_params), C.c_int32, C.c_int32, c_double_p, C.c_int64, C.c_uint64, C.c_int, C.c_int64,
      C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p]),
    ("lsmc_simulate_noalloc_max_periods", C.c_int32, []),
    ("lsmc_run", C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("lsmc_num_sim_periods", C.c_int32, [_H]),
    ("lsmc_num_periods", C.c_int32, [_H]),
    ("lsmc_get_grid", C.c_int32, [_H, C.c_int32, c_double_p, C.c_int32]),
    ("lsmc_num_evaluations", C.c_int64, [_H]),
    ("lsmc_get_results", C.c_int, [_H, C.POINTER(lsmc_results)]),
    ("lsmc_copy_panel", C.c_int, [_H, C.c_int, C.c_void_p]),
    ("lsmc_max_grid_points", C.c_int32, [_H]),
    ("lsmc_get_regression", C.c_int, [_H, c_double_p, c_double_p]),
    ("lsmc_last_error", C.c_char_p, [_H]),
    ("lsmc_destroy", None, [_H]),
    ("lsmc_comm_unique_id", C.c_int, [C.c_void_p]),
    ("lsmc_comm_init", C.c_int, [_H, C.c_void_p, C.c_int, C.c_int, C.c_int64]),
    ("lsmc_calculate_inventory_space", C.c_int, [C.POINTER(lsmc_storage), C.c_double, c_double_p, c_double_p]),
    ("lsmc_bang_bang_decision_set", C.c_int32,
     [C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, c_double_p]),
    ("lsmc_bisect_inventory_space", C.c_int, [c_double_p, C.c_int32, C.c_double, C.c_double, c_int32_p, c_int32_p]),
    ("lsmc_grid_points", C.c_int32, [C.c_double, C.c_double, C.c_double, c_double_p, C.c_int32]),
    ("lsmc_inject_withdraw_range", C.c_int, [C.POINTER(lsmc_storage), C.c_int32, C.c_double, c_double_p, c_double_p]),
    ("lsmc_inventory_space_bounds", C.c_int,
     [C.POINTER(lsmc_storage), C.c_int32, C.c_double, C.c_double, c_double_p, c_double_p]),
    ("lsmc_parse_basis", C.c_int32, [C.c_char_p, C.c_int32, c_int32_p, C.c_int32]),
]

_lib = None


def load_library():
    """Loads liblsmc_b200.so (built in-tree by __graft_entry__.build()). Raises if it is absent: there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "storage_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the library does not export what the header declares
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib
