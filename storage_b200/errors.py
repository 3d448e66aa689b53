"""Exceptions of the reference API surface and the status-code mapping of include/lsmc_b200.h."""
from . import _abi


class ArgumentException(ValueError):
    """System.ArgumentException (LsmcStorageValuation.cs:65,95,98; StorageHelper.cs:113,131,155,309,313)."""


class InventoryConstraintsCannotBeFulfilledException(Exception):
    """Exceptions/InventoryConstraintsCannotBeFulfilledException.cs (LsmcStorageValuation.cs:78; StorageHelper.cs:102)."""


class InvalidOperationException(RuntimeError):
    """System.InvalidOperationException raised by LsmcValuationParameters.Builder.Build (LsmcValuationParameters.cs:126-138)."""


class OperationCanceledException(Exception):
    """System.OperationCanceledException (LsmcStorageValuation.cs:345,521)."""


class CudaError(RuntimeError):
    """No B200 / CUDA failure. The product path never falls back to the CPU."""


class NumericalError(ArithmeticError):
    """Rank-deficient regression design matrix."""


class UnsupportedStorageFeature(NotImplementedError):
    """A reference overload that takes an opaque Func<> and so cannot be evaluated on the device."""


_STATUS_TO_EXC = {
    _abi.LSMC_ERR_ARGUMENT: ArgumentException,
    _abi.LSMC_ERR_INVENTORY_CONSTRAINTS: InventoryConstraintsCannotBeFulfilledException,
    _abi.LSMC_ERR_INVALID_OPERATION: InvalidOperationException,
    _abi.LSMC_ERR_CANCELLED: OperationCanceledException,
    _abi.LSMC_ERR_CUDA: CudaError,
    _abi.LSMC_ERR_NUMERICAL: NumericalError,
    _abi.LSMC_ERR_UNSUPPORTED: UnsupportedStorageFeature,
}


def check_status(status, handle):
    if status == _abi.LSMC_OK:
        return
    lib = _abi.load_library()
    msg = lib.lsmc_last_error(handle)
    raise _STATUS_TO_EXC.get(status, RuntimeError)((msg or b"").decode() or f"lsmc status {status}")
