"""Host-side mirror of the reference's LSMC API surface over the C ABI (include/lsmc_b200.h):

  LsmcValuationParameters<T>.Builder   LsmcValuation/LsmcValuationParameters.cs:85-255
  LsmcStorageValuation.Calculate       LsmcValuation/LsmcStorageValuation.cs:57-657
  LsmcStorageValuationResults<T>       LsmcValuation/LsmcStorageValuationResults.cs:34-80
  SimulationDataReturned               LsmcValuation/SimulationDataReturned.cs:31-47
  BasisFunctionsBuilder / Sim          LsmcValuation/BasisFunctions/*.cs
  FixedSpacingStateSpaceGridCalc       StateSpaceGridCalc/FixedSpacingStateSpaceGridCalc.cs:34-83
  StorageHelper                        StorageHelper.cs
  SpotSimResultsFromPanels             PythonHelpers/SpotSimResultsFromPanels.cs:36-117
  MultiFactorParameters                Cmdty.Core 0.2.0 (un-vendored; call sites LsmcValuationParameters.cs:153-196)

All numerics run in liblsmc_b200.so on the GPU; this module only resolves dates into per-period numbers, flattens the
builder objects into the lsmc_problem descriptor and marshals results.  There is no CPU path here.
"""
import ctypes as C
import enum
import math

import numpy as np

from . import _abi
from .errors import (ArgumentException, InvalidOperationException, InventoryConstraintsCannotBeFulfilledException,
                     check_status)
from .sharding import shard_bounds
from .time_period import Day


class SimulationDataReturned(enum.IntFlag):
    Nothing = 0  # `None` in C#
    SpotPricesForRegression = 1
    SpotPricesForValuation = 1 << 2
    AllSpotPrices = 1 | (1 << 2)
    FactorsForRegression = 1 << 3
    FactorsForValuation = 1 << 4
    AllFactors = (1 << 3) | (1 << 4)
    Inventory = 1 << 5
    InjectWithdrawVolume = 1 << 6
    CmdtyConsumed = 1 << 7
    InventoryLoss = 1 << 8
    NetVolume = 1 << 9
    Pv = 1 << 10
    All = 1 | (1 << 2) | (1 << 3) | (1 << 4) | (1 << 5) | (1 << 6) | (1 << 7) | (1 << 8) | (1 << 9) | (1 << 10)


# ----------------------------------------------------------------------------------------------------------------
# Basis functions (structured: every basis function is a power monomial, PowerMonomialBuilder.cs:33-34)
# ----------------------------------------------------------------------------------------------------------------
class PowerMonomialBuilder:
    def __init__(self, spotPower, markovFactorPowers=None):
        self.SpotPower = int(spotPower)
        self.MarkovFactorPowers = dict(markovFactorPowers or {})

    def __mul__(self, other):  # PowerMonomialBuilder.cs:44-57
        powers = dict(self.MarkovFactorPowers)
        for k, v in other.MarkovFactorPowers.items():
            powers[k] = powers.get(k, 0) + v
        return PowerMonomialBuilder(self.SpotPower + other.SpotPower, powers)

    def Pow(self, power):  # PowerMonomialBuilder.cs:59-65
        return PowerMonomialBuilder(self.SpotPower * power, {k: v * power for k, v in self.MarkovFactorPowers.items()})

    def __pow__(self, power):
        return self.Pow(power)

    def __add__(self, other):
        return BasisFunctionsBuilder([self]) + other

    def key(self):
        return (self.SpotPower, tuple(sorted((k, v) for k, v in self.MarkovFactorPowers.items() if v)))


class Sim:  # Sim.cs:30-46
    Spot = S = PowerMonomialBuilder(1)

    @staticmethod
    def Factor(i):
        return PowerMonomialBuilder(0, {i: 1})


for _i in range(10):
    setattr(Sim, f"X{_i}", Sim.Factor(_i))


class BasisFunctionsBuilder:
    """Ordered list of monomials with the `+` combinator (BasisFunctionsBuilder.cs:40-89)."""

    def __init__(self, functions):
        self.functions = list(functions)

    def __add__(self, other):
        if isinstance(other, PowerMonomialBuilder):
            return BasisFunctionsBuilder(self.functions + [other])
        return BasisFunctionsBuilder(self.functions + list(other.functions))

    def __iter__(self):
        return iter(self.functions)

    def __len__(self):
        return len(self.functions)

    Ones = None  # class attributes assigned below

    @staticmethod
    def SpotPricePower(power):
        return BasisFunctionsBuilder([PowerMonomialBuilder(power)])

    @staticmethod
    def MarkovFactorPower(markovFactor, power):
        return BasisFunctionsBuilder([PowerMonomialBuilder(0, {markovFactor: power})])

    @staticmethod
    def MarkovFactorAllPositiveIntegerPowersUpTo(markovFactor, maxPower):
        if maxPower < 1:
            raise ArgumentException("Maximum power must be greater than zero.")
        return BasisFunctionsBuilder([PowerMonomialBuilder(0, {markovFactor: i + 1}) for i in range(maxPower)])

    @staticmethod
    def AllMarkovFactorAllPositiveIntegerPowersUpTo(maxPower, numMarkovFactors):
        if maxPower < 1:
            raise ArgumentException("Maximum power must be greater than zero.")
        return BasisFunctionsBuilder([PowerMonomialBuilder(0, {f: i + 1}) for f in range(numMarkovFactors) for i in range(maxPower)])

    @staticmethod
    def Parse(expression, numFactors=_abi.LSMC_MAX_FACTORS):
        """BasisFunctionsBuilder.Parse (BasisFunctionsBuilder.cs:90-131) through the library's Roslyn-free parser."""
        lib = _abi.load_library()
        cap = _abi.LSMC_MAX_BASIS
        powers = (C.c_int32 * (cap * (1 + numFactors)))()
        k = lib.lsmc_parse_basis(expression.encode(), numFactors, powers, cap)
        if k < 0:
            check_status(-k, None)
        out = []
        for i in range(k):
            row = powers[i * (1 + numFactors):(i + 1) * (1 + numFactors)]
            out.append(PowerMonomialBuilder(row[0], {f: row[1 + f] for f in range(numFactors) if row[1 + f]}))
        return BasisFunctionsBuilder(out)


BasisFunctionsBuilder.Ones = BasisFunctionsBuilder([PowerMonomialBuilder(0)])


def basis_power_table(basis_functions, num_factors):
    fns = list(basis_functions)
    table = np.zeros((len(fns), 1 + num_factors), dtype=np.int32)
    for k, f in enumerate(fns):
        if f.SpotPower < 0 or any(v < 0 for v in f.MarkovFactorPowers.values()):
            raise ArgumentException("Spot power and Markov factor powers must be non-negative.")  # BasisFunctions.cs:95-98
        table[k, 0] = f.SpotPower
        for i, v in f.MarkovFactorPowers.items():
            if v and i >= num_factors:
                raise ArgumentException(f"Basis function uses factor {i} but the simulation has {num_factors} factors.")
            if i < num_factors:
                table[k, 1 + i] = v
    return table


# ----------------------------------------------------------------------------------------------------------------
# Grid calc, discounting helpers
# ----------------------------------------------------------------------------------------------------------------
class FixedSpacingStateSpaceGridCalc:
    def __init__(self, spacing):
        if spacing <= 0.0:
            raise ArgumentException("Parameter must be positive")
        self.Spacing = float(spacing)

    def GetGridPoints(self, lower, upper):
        lib = _abi.load_library()
        n = lib.lsmc_grid_points(lower, upper, self.Spacing, None, 0)
        if n < 0:
            check_status(-n, None)
        out = np.zeros(n, dtype=np.float64)
        lib.lsmc_grid_points(lower, upper, self.Spacing, out.ctypes.data_as(_abi.c_double_p), n)
        return out

    @staticmethod
    def CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, numGridPointsOverGlobalInventoryRange):
        if numGridPointsOverGlobalInventoryRange < 3:  # FixedSpacingStateSpaceGridCalc.cs:71-72
            raise ArgumentException("numGridPointsOverGlobalInventoryRange value must be at least 3.")
        periods = storage.StartPeriod.EnumerateTo(storage.EndPeriod)
        gmax = max(storage.MaxInventory(p) for p in periods)
        gmin = min(storage.MinInventory(p) for p in periods)
        return FixedSpacingStateSpaceGridCalc((gmax - gmin) / (numGridPointsOverGlobalInventoryRange - 1))


class StorageHelper:
    @staticmethod
    def CreateAct65ContCompDiscounter(rate):
        """StorageHelper.cs:262-276: exp(-days/365 * r), 1 for cash flows on or before the present day."""
        rate_fn = rate if callable(rate) else (lambda d: rate)

        def discount(presentDay, cashFlowDay):
            if cashFlowDay <= presentDay:
                return 1.0
            return math.exp(-cashFlowDay.OffsetFrom(presentDay) / 365.0 * rate_fn(cashFlowDay))
        return discount

    @staticmethod
    def CalculateBangBangDecisionSet(injectWithdrawRange, currentInventory, inventoryLoss, nextStepMinInventory,
                                     nextStepMaxInventory, numericalTolerance, numExtraDecisions=0):
        lib = _abi.load_library()
        out = (C.c_double * _abi.LSMC_MAX_DECISIONS)()
        n = lib.lsmc_bang_bang_decision_set(injectWithdrawRange.MinInjectWithdrawRate, injectWithdrawRange.MaxInjectWithdrawRate,
                                            currentInventory, inventoryLoss, nextStepMinInventory, nextStepMaxInventory,
                                            numericalTolerance, numExtraDecisions, out)
        if n < 0:
            check_status(-n, None)
        return list(out[:n])

    @staticmethod
    def BisectInventorySpace(inventoryGrid, inventory, numericalTolerance):
        lib = _abi.load_library()
        g = np.ascontiguousarray(inventoryGrid, dtype=np.float64)
        lo, hi = C.c_int32(), C.c_int32()
        check_status(lib.lsmc_bisect_inventory_space(g.ctypes.data_as(_abi.c_double_p), len(g), inventory, numericalTolerance,
                                                     C.byref(lo), C.byref(hi)), None)
        return lo.value, hi.value

    @staticmethod
    def CalculateInventorySpace(storage, startingInventory, currentPeriod):
        if currentPeriod > storage.EndPeriod:
            raise ArgumentException("Storage has expired")
        start_active = max(storage.StartPeriod, currentPeriod)
        flat = storage.flatten(start_active, lambda d: 1.0)
        P = flat.num_periods
        mn, mx = np.zeros(max(P, 1)), np.zeros(max(P, 1))
        lib = _abi.load_library()
        check_status(lib.lsmc_calculate_inventory_space(C.byref(flat.struct), float(startingInventory),
                                                        mn.ctypes.data_as(_abi.c_double_p), mx.ctypes.data_as(_abi.c_double_p)), None)
        return {start_active.Offset(i + 1): (mn[i], mx[i]) for i in range(P)}


# ----------------------------------------------------------------------------------------------------------------
# Multi-factor model parameters (Cmdty.Core MultiFactorParameters; call sites LVP.cs:153-196)
# ----------------------------------------------------------------------------------------------------------------
class MultiFactorParameters:
    """factors: list of (mean_reversion, vol) with vol a float or a callable period -> vol; corr: FxF matrix."""

    def __init__(self, factors, factor_corrs):
        self.factors = [(float(mr), vol) for mr, vol in factors]
        self.corr = np.array(factor_corrs, dtype=np.float64)
        n = len(self.factors)
        if self.corr.shape != (n, n):
            raise ArgumentException("factor correlation matrix has the wrong shape")

    @property
    def NumFactors(self):
        return len(self.factors)

    @staticmethod
    def For1Factor(meanReversion, spotVol):
        return MultiFactorParameters([(meanReversion, spotVol)], [[1.0]])

    @staticmethod
    def For2Factors(correlation, factor1, factor2):
        return MultiFactorParameters([factor1, factor2], [[1.0, correlation], [correlation, 1.0]])

    @staticmethod
    def For3FactorSeasonal(spotMeanReversion, spotVol, longTermVol, seasonalVol, start, end):
        """Factor order spot, long-term, seasonal; identity correlation; seasonal vol curve
        0.5*seasonalVol*sin(2*pi*t_from_1Feb + pi/2) with a 365.25-day year
        (src/Cmdty.Storage.Python/cmdty_storage/multi_factor_diffusion_model.py:137-172)."""
        peak = Day(start.Start.year, 2, 1)

        def seasonal(p):
            t = (p.First().ordinal - peak.ordinal) / 365.25
            return math.sin(2.0 * math.pi * t + math.pi / 2.0) * (seasonalVol / 2.0)
        return MultiFactorParameters([(spotMeanReversion, spotVol), (0.0, longTermVol), (0.0, seasonal)], np.eye(3))

    def vol(self, i, period):
        v = self.factors[i][1]
        if callable(v):
            return float(v(period))
        if isinstance(v, dict):
            return float(v[period])
        return float(v)


class SpotSimResultsFromPanels:
    """Caller-supplied simulations (SpotSimResultsFromPanels.cs:44-74): spot [rows][N] + one [rows][N] panel per factor,
    rows keyed by period."""

    def __init__(self, periods, spotPriceSims, factorSims):
        spot = np.ascontiguousarray(spotPriceSims, dtype=np.float64)
        if spot.ndim != 2 or spot.shape[1] == 0:
            raise ArgumentException("spotPriceSims.NumCols (number of simulations), cannot equal zero.")
        facs = [np.ascontiguousarray(f, dtype=np.float64) for f in factorSims]
        for i, f in enumerate(facs):
            if f.shape[1:] != spot.shape[1:]:
                raise ArgumentException(f"Inconsistent number of columns (simulations) between spotPriceSims and item {i} of factorSims.")
            if f.shape[0] != spot.shape[0]:
                raise ArgumentException(f"Inconsistent number of rows (time periods) between spotPriceSims and item {i} of factorSims.")
        self.SimulatedPeriods = list(periods)
        if len(self.SimulatedPeriods) != spot.shape[0]:
            raise ArgumentException("Number of row keys differs from the number of rows.")
        self.SpotPriceSims = spot
        self.FactorSims = facs

    @property
    def NumSims(self):
        return self.SpotPriceSims.shape[1]

    @property
    def NumFactors(self):
        return len(self.FactorSims)

    def rows_for(self, periods):
        """Rows of the panels that hold `periods`: a slice when they are consecutive (the usual case: no copy of the panels is
        made on the way to the device), else an index list."""
        index = {p: i for i, p in enumerate(self.SimulatedPeriods)}
        rows = [index[p] for p in periods]
        if rows and rows == list(range(rows[0], rows[0] + len(rows))):
            return slice(rows[0], rows[0] + len(rows))
        return rows


# ----------------------------------------------------------------------------------------------------------------
# Parameters
# ----------------------------------------------------------------------------------------------------------------
class LsmcValuationParameters:
    def __init__(self, b):
        self.CurrentPeriod = b.CurrentPeriod
        self.Inventory = float(b.Inventory)
        self.ForwardCurve = b.ForwardCurve
        self.Storage = b.Storage
        self.SettleDateRule = b.SettleDateRule
        self.DiscountFactors = b.DiscountFactors
        self.GridCalc = b.GridCalc
        self.NumericalTolerance = b.NumericalTolerance
        self.BasisFunctions = list(b.BasisFunctions)
        self.CancellationToken = b.CancellationToken
        self.OnProgressUpdate = b.OnProgressUpdate
        self.DiscountDeltas = bool(b.DiscountDeltas)
        self.ExtraDecisions = int(b.ExtraDecisions)
        self.SimulationDataReturned = SimulationDataReturned(b.SimulationDataReturned)
        self._sim_spec = b._sim_spec
        self.ReferenceCompatTerminalPv = bool(b.ReferenceCompatTerminalPv)

    class Builder:
        DefaultNumericalTolerance = 1e-10  # LVP.cs:88

        def __init__(self):
            self.Inventory = None
            self.ForwardCurve = None
            self.Storage = None
            self.SettleDateRule = None
            self.DiscountFactors = None
            self.GridCalc = None
            self.NumericalTolerance = self.DefaultNumericalTolerance
            self.BasisFunctions = None
            self.CancellationToken = None
            self.OnProgressUpdate = None
            self.ExtraDecisions = 0
            self.SimulationDataReturned = SimulationDataReturned.Nothing
            self.DiscountDeltas = False
            self.ReferenceCompatTerminalPv = True
            self._current = None
            self._current_set = False
            self._sim_spec = None

        @property
        def CurrentPeriod(self):
            return self._current

        @CurrentPeriod.setter
        def CurrentPeriod(self, value):
            self._current_set = True
            self._current = value

        def Build(self):  # LVP.cs:124-144
            if not self._current_set:
                raise InvalidOperationException("CurrentPeriod has not been set.")
            for name in ("Inventory", "ForwardCurve", "Storage", "SettleDateRule", "DiscountFactors", "GridCalc"):
                if getattr(self, name) is None:
                    raise InvalidOperationException(name + " has not been set.")
            if self._sim_spec is None:
                raise InvalidOperationException("RegressionSpotSimsGenerator has not been set.")
            if self.BasisFunctions is None:
                raise InvalidOperationException("BasisFunctions has not been set.")
            if self.ExtraDecisions < 0:
                raise InvalidOperationException("ExtraDecisions must be non-negative.")
            return LsmcValuationParameters(self)

        def SimulateWithMultiFactorModelAndMersenneTwister(self, modelParameters, numSims, simSeed=None, valuationSimSeed=None,
                                                           antithetic=True):
            """LVP.cs:185-196.  The device simulator is counter-based (Philox), not a Mersenne Twister: draws differ from
            the reference's stream, moments agree (north star).  valuationSimSeed None continues the regression stream."""
            if modelParameters is None:
                raise ArgumentException("modelParameters is null")
            if numSims <= 0:
                raise ArgumentException("Number of simulations must be positive.")
            seed = 0x5EED if simSeed is None else int(simSeed)
            val_seed = (seed ^ 0x9E3779B97F4A7C15) if valuationSimSeed is None else int(valuationSimSeed)
            self._sim_spec = ("model", modelParameters, int(numSims), seed & (2**64 - 1), val_seed & (2**64 - 1), bool(antithetic))
            return self

        def UseSpotSimResults(self, regressionSpotSim, valuationSpotSim):  # LVP.cs:198-216
            if regressionSpotSim is None:
                raise ArgumentException("regressionSpotSim is null")
            if valuationSpotSim is None:
                raise ArgumentException("valuationSpotSim is null")
            self._sim_spec = ("panels", regressionSpotSim, valuationSpotSim)
            return self

        def Clone(self):  # LVP.cs:233-253
            c = LsmcValuationParameters.Builder()
            c.__dict__.update(self.__dict__)
            if self.BasisFunctions is not None:
                c.BasisFunctions = list(self.BasisFunctions)
            return c


class FlatProblem:
    """lsmc_problem descriptor + the numpy arrays that back its pointers."""

    def __init__(self, params, num_sims, num_factors, extra_flags=0):
        st = params.Storage
        cur = params.CurrentPeriod
        start_active = max(st.StartPeriod, cur)
        present = cur.First()

        def df(day):
            return params.DiscountFactors(present, day)
        self.flat_storage = st.flatten(start_active, df)
        P = self.flat_storage.num_periods
        periods = self.flat_storage.periods
        self.periods = periods
        self.start_active = start_active
        # forward-curve coverage (LSV.cs:94-98)
        keys = sorted(params.ForwardCurve)
        if keys[0] > start_active:
            raise ArgumentException(f"Forward curve starts too late. Must start on or before the period {start_active}.")
        if keys[-1] < st.EndPeriod:
            raise ArgumentException("Forward curve does not extend until storage end period.")
        self.forward = np.array([params.ForwardCurve[p] for p in periods], dtype=np.float64)
        self.settle_df = np.array([df(params.SettleDateRule(p)) for p in periods[:P]] or [1.0], dtype=np.float64)
        self.powers = basis_power_table(params.BasisFunctions, num_factors)
        pb = _abi.lsmc_problem()
        pb.storage = self.flat_storage.struct
        pb.inventory = params.Inventory
        pb.current_period_is_active = 1 if cur >= st.StartPeriod else 0
        pb.forward_price = self.forward.ctypes.data_as(_abi.c_double_p)
        pb.settle_df = self.settle_df.ctypes.data_as(_abi.c_double_p)
        pb.grid_spacing = params.GridCalc.Spacing
        pb.numerical_tolerance = params.NumericalTolerance
        pb.extra_decisions = params.ExtraDecisions
        pb.discount_deltas = 1 if params.DiscountDeltas else 0
        pb.simulation_data_returned = int(params.SimulationDataReturned) | int(extra_flags)
        pb.num_factors = num_factors
        pb.num_basis = self.powers.shape[0]
        pb.basis_powers = self.powers.ctypes.data_as(_abi.c_int32_p)
        pb.num_sims = num_sims
        pb.reference_compat_terminal_pv = 1 if params.ReferenceCompatTerminalPv else 0
        self.struct = pb
        self.num_periods = P
        self.first_sim = pb.current_period_is_active
        self.sim_periods = periods[self.first_sim:]

    def signature(self):
        """Hashable digest of everything lsmc_create reads: two problems with the same signature build the same tables."""
        import hashlib
        pb = self.struct
        m = hashlib.blake2b(digest_size=16)
        for arr in list(self.flat_storage.arrays.values()) + [self.forward, self.settle_df, self.powers]:
            m.update(np.ascontiguousarray(arr).tobytes())
        st = self.flat_storage.struct
        scalars = (st.num_periods, st.must_be_empty_at_end, st.num_ratchet_rows, tuple(st.terminal_npv_coeff), st.newton_raphson_accuracy,
                   st.newton_raphson_max_iterations, st.newton_raphson_subdivision,
                   pb.inventory, pb.current_period_is_active, pb.grid_spacing, pb.numerical_tolerance, pb.extra_decisions, pb.discount_deltas,
                   pb.simulation_data_returned, pb.num_factors, pb.num_basis, pb.num_sims, pb.reference_compat_terminal_pv)
        return (m.hexdigest(),) + scalars

    def model_arrays(self, model, current):
        """lsmc_multi_factor_params for periods 0..P (vol of the contract delivering in each period, Act/365 times)."""
        F = model.NumFactors
        P = self.num_periods
        vol = np.array([[model.vol(i, p) for p in self.periods] for i in range(F)], dtype=np.float64)
        mr = np.array([f[0] for f in model.factors], dtype=np.float64)
        corr = np.ascontiguousarray(model.corr, dtype=np.float64)
        t = np.array([(p.First().ordinal - current.First().ordinal) / 365.0 for p in self.periods], dtype=np.float64)
        mp = _abi.lsmc_multi_factor_params()
        mp.num_factors = F
        mp.mean_reversion = mr.ctypes.data_as(_abi.c_double_p)
        mp.factor_vol = vol.ctypes.data_as(_abi.c_double_p)
        mp.factor_corr = corr.ctypes.data_as(_abi.c_double_p)
        mp.period_time = t.ctypes.data_as(_abi.c_double_p)
        assert vol.shape == (F, P + 1)
        return mp, (vol, mr, corr, t)


# ----------------------------------------------------------------------------------------------------------------
# Results
# ----------------------------------------------------------------------------------------------------------------
class StorageProfile:
    def __init__(self, inventory, injectWithdrawVolume, cmdtyConsumed, inventoryLoss, periodPv):
        self.Inventory = inventory
        self.InjectWithdrawVolume = injectWithdrawVolume
        self.CmdtyConsumed = cmdtyConsumed
        self.InventoryLoss = inventoryLoss
        self.PeriodPv = periodPv

    @property
    def NetVolume(self):
        return -self.InjectWithdrawVolume - self.CmdtyConsumed


class TriggerPrices:
    def __init__(self, t):
        self.MaxInjectVolume = t.max_inject_volume if t.has_inject_price else None
        self.MaxInjectTriggerPrice = t.max_inject_trigger_price if t.has_inject_price else None
        self.MaxWithdrawVolume = t.max_withdraw_volume if t.has_withdraw_price else None
        self.MaxWithdrawTriggerPrice = t.max_withdraw_trigger_price if t.has_withdraw_price else None
        self.HasInjectPrice = bool(t.has_inject_price)
        self.HasWithdrawPrice = bool(t.has_withdraw_price)


class TriggerPriceVolumeProfiles:
    def __init__(self, t):
        self.InjectTriggerPrices = [(t.inject_volume[i], t.inject_price[i]) for i in range(t.num_inject)]
        self.WithdrawTriggerPrices = [(t.withdraw_volume[i], t.withdraw_price[i]) for i in range(t.num_withdraw)]


class LsmcStorageValuationResults:
    """LsmcStorageValuationResults.cs:37-53; time series are dicts keyed by period, panels are numpy [rows][N]."""

    def __init__(self):
        self.Npv = 0.0
        self.ValuationSimStandardError = 0.0
        self.Deltas = {}
        self.ExpectedStorageProfile = {}
        self.RegressionSpotPriceSim = np.zeros((0, 0))
        self.ValuationSpotPriceSim = np.zeros((0, 0))
        self.InventoryBySim = np.zeros((0, 0))
        self.InjectWithdrawVolumeBySim = np.zeros((0, 0))
        self.CmdtyConsumedBySim = np.zeros((0, 0))
        self.InventoryLossBySim = np.zeros((0, 0))
        self.NetVolumeBySim = np.zeros((0, 0))
        self.PvByPeriodAndSim = np.zeros((0, 0))
        self.PvBySim = np.zeros(0)
        self.TriggerPriceVolumeProfiles = {}
        self.TriggerPrices = {}
        self.RegressionMarkovFactors = []
        self.ValuationMarkovFactors = []
        # extras without a reference counterpart
        self.BackwardNpv = 0.0
        self.Timings = {}
        self.KernelLaunches = 0
        self.DecisionIndexBySim = None

    @staticmethod
    def CreateExpiredResults():
        return LsmcStorageValuationResults()

    @staticmethod
    def CreateEndPeriodResults(npv):
        r = LsmcStorageValuationResults()
        r.Npv = npv
        return r


class Handle:
    """RAII wrapper of lsmc_handle."""

    def __init__(self, flat, device=0):
        self.lib = _abi.load_library()
        self.h = C.c_void_p()
        self.flat = flat
        check_status(self.lib.lsmc_create(C.byref(flat.struct), device, C.byref(self.h)), None)

    def check(self, status):
        check_status(status, self.h)

    def close(self):
        if self.h:
            self.lib.lsmc_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_sims(self, which, spot, factors, wait=True):
        """wait=False: lsmc_set_sims_async; the returned arrays must be kept alive until lsmc_run has returned."""
        spot = np.ascontiguousarray(spot, dtype=np.float64)
        facs = [np.ascontiguousarray(f, dtype=np.float64) for f in factors]
        ptrs = (C.c_void_p * max(len(facs), 1))(*[f.ctypes.data for f in facs])
        fn = self.lib.lsmc_set_sims if wait else self.lib.lsmc_set_sims_async
        self.check(fn(self.h, which, spot.ctypes.data, ptrs))
        return spot, facs

    def panel(self, which, rows, n, lead=None):
        shape = (rows, n) if lead is None else (lead, rows, n)
        out = np.zeros(shape, dtype=np.float64)
        self.check(self.lib.lsmc_copy_panel(self.h, which, out.ctypes.data))
        return out


class PathShard:
    """This process's share of a valuation whose paths are sharded over several GPUs, one process per GPU (no reference
    counterpart).  comm_id is the LSMC_COMM_ID_BYTES blob rank 0 drew with lsmc_comm_unique_id and sent to the
    other ranks by whatever channel the host has (bench.py: torch.distributed broadcast)."""

    def __init__(self, rank, world, comm_id, total_sims):
        self.rank, self.world, self.comm_id, self.total_sims = int(rank), int(world), bytes(comm_id), int(total_sims)
        self.begin, self.end = shard_bounds(self.total_sims, self.world, self.rank)

    @property
    def num_sims(self):
        return self.end - self.begin


class LsmcStorageValuation:
    WithNoLogger = None  # assigned below

    def __init__(self, logger=None, device=0, shard=None, reuse_handle=False):
        """shard: PathShard of this process (None: all paths on this GPU).  reuse_handle: keep the device-side state of the
        last valuation (buffers, tables) and run the next one on it when it has the same problem descriptor; a host that
        revalues the same storage repeatedly (new simulations, same curves) then pays no table build or device allocation.
        Call Close() to release it."""
        self._logger = logger
        self._device = device
        self._shard = shard
        self._reuse = bool(reuse_handle)
        self._kept = None  # (signature, Handle)

    def Close(self):
        if self._kept is not None:
            self._kept[1].close()
            self._kept = None

    def __del__(self):
        try:
            self.Close()
        except Exception:
            pass

    def _handle_for(self, flat):
        if not self._reuse:
            return self._open(flat)
        sig = flat.signature() + (self._device, None if self._shard is None else (self._shard.rank, self._shard.world, self._shard.comm_id))
        if self._kept is not None and self._kept[0] == sig:
            self._kept[1].flat = flat
            return self._kept[1]
        self.Close()
        h = self._open(flat)
        self._kept = (sig, h)
        return h

    def _open(self, flat):
        h = Handle(flat, self._device)
        if self._shard is not None and self._shard.world > 1:
            s = self._shard
            h.check(h.lib.lsmc_comm_init(h.h, s.comm_id, s.rank, s.world, s.total_sims))
        return h

    # per-simulation result panels larger than this are streamed to the host during the run instead of being kept on the device
    STREAM_PANELS_ABOVE_BYTES = 1 << 30

    def Calculate(self, lsmcParams, keep_decisions=False, reference_operation_order=False, stream_panels=None):
        """LsmcStorageValuation.Calculate (LSV.cs:57-657).

        keep_decisions, reference_operation_order and stream_panels have no reference counterpart: the first keeps the argmax
        index panel, the second evaluates each decision's cash flow in the reference's own sequence of roundings
        (LSV.cs:317-336) instead of the default fused form (include/lsmc_b200.h LSMC_REFERENCE_OPERATION_ORDER), the third
        forces (True) or forbids (False) delivering the requested per-simulation panels block by block while the forward
        simulation runs (LSMC_STREAM_PANELS; default: when they exceed STREAM_PANELS_ABOVE_BYTES)."""
        p = lsmcParams
        notify = p.OnProgressUpdate
        if p.Inventory < 0:  # LSV.cs:64-65
            raise ArgumentException("Inventory cannot be negative.")
        if p.CurrentPeriod > p.Storage.EndPeriod:  # LSV.cs:67-71
            if notify:
                notify(1.0)
            return LsmcStorageValuationResults.CreateExpiredResults()
        if p.CurrentPeriod == p.Storage.EndPeriod:  # LSV.cs:73-87
            if p.Storage.MustBeEmptyAtEnd:
                if p.Inventory > 0:
                    raise InventoryConstraintsCannotBeFulfilledException("Storage must be empty at end, but inventory is greater than zero.")
                if notify:
                    notify(1.0)
                return LsmcStorageValuationResults.CreateExpiredResults()
            npv = p.Storage.TerminalStorageNpv(p.ForwardCurve[p.CurrentPeriod], p.Inventory)
            if notify:
                notify(1.0)
            return LsmcStorageValuationResults.CreateEndPeriodResults(npv)

        spec = p._sim_spec
        shard = self._shard
        first_path = 0
        if spec[0] == "panels":  # sharded: the panels hold this rank's slice of the paths
            reg, val = spec[1], spec[2]
            num_sims, num_factors = reg.NumSims, reg.NumFactors
            if shard is not None and num_sims != shard.num_sims:
                raise ArgumentException(f"the simulation panels hold {num_sims} paths, this rank's share is {shard.num_sims}")
        else:
            model, num_sims = spec[1], spec[2]
            num_factors = model.NumFactors
            if shard is not None:
                if num_sims != shard.total_sims:
                    raise ArgumentException("PathShard.total_sims differs from the number of simulations of the parameters")
                num_sims, first_path = shard.num_sims, shard.begin
        S = SimulationDataReturned
        result_flags = [S.Inventory, S.InjectWithdrawVolume, S.CmdtyConsumed, S.InventoryLoss, S.NetVolume, S.Pv]
        num_result_panels = sum(1 for f in result_flags if p.SimulationDataReturned & f) + (1 if keep_decisions else 0)
        if stream_panels is None:
            rows = p.Storage.EndPeriod.OffsetFrom(max(p.Storage.StartPeriod, p.CurrentPeriod)) + 1
            stream_panels = num_result_panels * rows * num_sims * 8 > self.STREAM_PANELS_ABOVE_BYTES
        stream_panels = bool(stream_panels) and num_result_panels > 0
        flat = FlatProblem(p, num_sims, num_factors, extra_flags=(_abi.LSMC_RETURN_DECISION_INDEX if keep_decisions else 0)
                           | (_abi.LSMC_REFERENCE_OPERATION_ORDER if reference_operation_order else 0)
                           | (_abi.LSMC_STREAM_PANELS if stream_panels else 0))
        if spec[0] == "panels":  # CheckSpotSim (LVP.cs:219-231): host-side, before any device call
            for name, sims in (("regressionSpotSim", reg), ("valuationSpotSim", val)):
                have = set(sims.SimulatedPeriods)
                for sp in flat.sim_periods:
                    if sp not in have:
                        raise ArgumentException(f"{name} does not contain simulations for period {sp}.")
        h = self._handle_for(flat)
        try:
            lib = h.lib
            keep_alive = []
            if spec[0] == "panels":
                # the regression set is needed by the first kernel; the valuation set only by the forward simulation, so it is
                # handed over without waiting and travels while the backward induction runs
                for sims, which in ((reg, 0), (val, 1)):
                    rows = sims.rows_for(flat.sim_periods)
                    keep_alive.append(h.set_sims(which, sims.SpotPriceSims[rows], [f[rows] for f in sims.FactorSims], wait=(which == 0)))
            else:
                _, model, _, seed, val_seed, antithetic = spec
                mp, keep = flat.model_arrays(model, p.CurrentPeriod)
                h.check(lib.lsmc_simulate(h.h, 0, C.byref(mp), seed, int(antithetic), first_path))
                h.check(lib.lsmc_simulate(h.h, 1, C.byref(mp), val_seed, int(antithetic), first_path))
                del keep
            cancel = C.c_int(0)
            token = p.CancellationToken

            def on_progress(value, _user):
                if notify:
                    notify(value)
                if token is not None and token.is_set():
                    cancel.value = 1
            cb = _abi.PROGRESS_FN(on_progress)
            if token is not None and token.is_set():
                cancel.value = 1
            streamed = {}
            if stream_panels:
                P = flat.num_periods
                for flag, which in zip(result_flags + [None], (_abi.LSMC_PANEL_INVENTORY, _abi.LSMC_PANEL_INJECT_WITHDRAW_VOLUME,
                                                               _abi.LSMC_PANEL_CMDTY_CONSUMED, _abi.LSMC_PANEL_INVENTORY_LOSS,
                                                               _abi.LSMC_PANEL_NET_VOLUME, _abi.LSMC_PANEL_PV, _abi.LSMC_PANEL_DECISION_INDEX)):
                    wanted = keep_decisions if flag is None else bool(p.SimulationDataReturned & flag)
                    if wanted:
                        streamed[which] = np.empty((P if flag is None else P + 1, num_sims), dtype=np.float64)
                        h.check(lib.lsmc_set_panel_sink(h.h, which, streamed[which].ctypes.data))
            h.check(lib.lsmc_run(h.h, C.cast(cb, C.c_void_p), None, C.addressof(cancel)))
            del keep_alive
            return self._collect(h, flat, p, num_sims, num_factors, keep_decisions, streamed)
        finally:
            if not self._reuse:
                h.close()

    @staticmethod
    def _collect(h, flat, p, num_sims, num_factors, keep_decisions, streamed=None):
        P = flat.num_periods
        periods = flat.periods
        deltas = np.zeros(P + 1)
        profile = np.zeros((P + 1, 5))
        trig = (_abi.lsmc_trigger_prices * max(P, 1))()
        pv = np.zeros(num_sims)
        r = _abi.lsmc_results()
        r.deltas = deltas.ctypes.data_as(_abi.c_double_p)
        r.profile = profile.ctypes.data_as(_abi.c_double_p)
        r.trigger_prices = trig
        r.pv_by_sim = pv.ctypes.data_as(_abi.c_double_p)
        h.check(h.lib.lsmc_get_results(h.h, C.byref(r)))
        out = LsmcStorageValuationResults()
        out.Npv = r.npv
        out.ValuationSimStandardError = r.std_error
        out.BackwardNpv = r.backward_npv
        out.Deltas = {periods[i]: deltas[i] for i in range(P + 1)}
        out.ExpectedStorageProfile = {periods[i]: StorageProfile(*profile[i]) for i in range(P + 1)}
        out.TriggerPrices = {periods[i]: TriggerPrices(trig[i]) for i in range(P)}
        out.TriggerPriceVolumeProfiles = {periods[i]: TriggerPriceVolumeProfiles(trig[i]) for i in range(P)}
        out.PvBySim = pv
        out.Timings = {k: getattr(r, "ms_" + k) for k in ("total", "regression_sim", "valuation_sim", "backward", "pseudo_inverse", "forward")}
        out.KernelLaunches = r.kernel_launches
        out.Timings["backward_kernel"] = r.ms_backward_kernel
        out.Timings["device"] = r.ms_device
        out.Timings["upload"] = r.ms_upload
        out.BackwardKernelLaunches = r.backward_kernel_launches
        out.SnappedInventories = r.snapped_inventories
        flags = p.SimulationDataReturned
        S = SimulationDataReturned
        rows = len(flat.sim_periods)
        streamed = streamed or {}

        def result_panel(which, nrows):  # streamed panels were delivered into their sink arrays during the run
            return streamed[which] if which in streamed else h.panel(which, nrows, num_sims)
        if flags & S.Inventory:
            out.InventoryBySim = result_panel(_abi.LSMC_PANEL_INVENTORY, P + 1)
        if flags & S.InjectWithdrawVolume:
            out.InjectWithdrawVolumeBySim = result_panel(_abi.LSMC_PANEL_INJECT_WITHDRAW_VOLUME, P + 1)
        if flags & S.CmdtyConsumed:
            out.CmdtyConsumedBySim = result_panel(_abi.LSMC_PANEL_CMDTY_CONSUMED, P + 1)
        if flags & S.InventoryLoss:
            out.InventoryLossBySim = result_panel(_abi.LSMC_PANEL_INVENTORY_LOSS, P + 1)
        if flags & S.NetVolume:
            out.NetVolumeBySim = result_panel(_abi.LSMC_PANEL_NET_VOLUME, P + 1)
        if flags & S.Pv:
            out.PvByPeriodAndSim = result_panel(_abi.LSMC_PANEL_PV, P + 1)
        if flags & S.SpotPricesForRegression:
            out.RegressionSpotPriceSim = h.panel(_abi.LSMC_PANEL_SPOT_REGRESSION, rows, num_sims)
        if flags & S.SpotPricesForValuation:
            out.ValuationSpotPriceSim = h.panel(_abi.LSMC_PANEL_SPOT_VALUATION, rows, num_sims)
        if flags & S.FactorsForRegression:
            out.RegressionMarkovFactors = list(h.panel(_abi.LSMC_PANEL_FACTORS_REGRESSION, rows, num_sims, num_factors))
        if flags & S.FactorsForValuation:
            out.ValuationMarkovFactors = list(h.panel(_abi.LSMC_PANEL_FACTORS_VALUATION, rows, num_sims, num_factors))
        if keep_decisions:
            out.DecisionIndexBySim = result_panel(_abi.LSMC_PANEL_DECISION_INDEX, P)
        out.SimulatedPeriods = list(flat.sim_periods)
        return out


class _NoLogger:
    def __get__(self, obj, objtype=None):
        return LsmcStorageValuation()


LsmcStorageValuation.WithNoLogger = _NoLogger()
