"""storage_b200 — B200-native LSMC storage valuation: the host-side mirror of cmdty/storage's API for the
LsmcStorageValuation.Calculate path, over the C ABI of liblsmc_b200.so (include/lsmc_b200.h)."""
from .errors import (ArgumentException, CudaError, InvalidOperationException,  # noqa: F401
                     InventoryConstraintsCannotBeFulfilledException, NumericalError, OperationCanceledException,
                     UnsupportedStorageFeature)
from .time_period import Day, Month, settle_on_20th_of_following_month  # noqa: F401
from .cmdty_storage import (CmdtyStorage, ConstantInjectWithdrawConstraint, InjectWithdrawRange,  # noqa: F401
                            PiecewiseLinearInjectWithdrawConstraint, PolynomialInjectWithdrawConstraint,
                            StepInjectWithdrawConstraint)
from .lsmc import (BasisFunctionsBuilder, FixedSpacingStateSpaceGridCalc, FlatProblem, Handle,  # noqa: F401
                   LsmcStorageValuation, LsmcStorageValuationResults, LsmcValuationParameters, MultiFactorParameters, PathShard,
                   PowerMonomialBuilder, Sim, SimulationDataReturned, SpotSimResultsFromPanels, StorageHelper)
