"""The C-ABI library loads without a GPU and exports every symbol include/lsmc_b200.h declares; compute entry points fail
loudly (LSMC_ERR_CUDA) instead of falling back to the CPU."""
import ctypes as C
import os
import re

import pytest

from storage_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "lsmc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lsmc_[a-z_0-9]+)\s*\(", text)) - {"lsmc_progress_fn"})


def test_library_exports_every_declared_symbol():
    lib = _abi.load_library()
    declared = header_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/lsmc_b200.h but not exported"
    bound = {n for n, _, _ in _abi.SYMBOLS}
    assert bound == set(declared), (bound ^ set(declared))


def test_ctypes_struct_sizes_match_the_header_layout():
    assert C.sizeof(_abi.lsmc_trigger_prices) == 4 * 2 + 8 * 10 * 4 + 4 * 2 + 8 * 4
    assert C.sizeof(_abi.lsmc_storage) % 8 == 0 and C.sizeof(_abi.lsmc_problem) % 8 == 0


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from storage_b200 import CudaError
    from tests import cases
    with pytest.raises(CudaError):
        cases.run_gpu(cases.simple_one_factor(num_sims=64, num_days=5, inventory=1000.0))


def test_product_does_not_import_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "storage_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".hpp", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower() or f == "_abi.py" and "oracle" not in text.lower(), f"{f} mentions the oracle"
