"""The CUDA library against numbers the REFERENCE produced (north star: "results must match the reference C# Lsmc on identical
pre-simulated spot paths, fed through its caller-supplied-simulation overload: NPV, per-step decisions and deltas within 1e-9
relative").  The paths are the reference's own (oracle/reference_sim.py reproduces its simulator's golden values digit for
digit, tests/test_reference_pinning.py), fed through UseSpotSimResults; the expected values are the literals and CSV fixtures of
the reference's regression tests (src/Cmdty.Storage.Python/tests/test_multi_factor.py:99-124, 259-270)."""
import numpy as np
import pytest

from storage_b200 import LsmcStorageValuation, SimulationDataReturned
from tests import reference_goldens

pytestmark = pytest.mark.gpu

REL = 1e-9


@pytest.mark.parametrize("name", ["multi_factor_test-1", "three_factor_test-1"])
@pytest.mark.parametrize("form", ["fused", "reference-order"])
def test_cuda_valuation_equals_the_reference_goldens(name, form):
    case = reference_goldens.build(name, flags=SimulationDataReturned.All)
    notified = []
    case.params.OnProgressUpdate = notified.append
    res = LsmcStorageValuation().Calculate(case.params, reference_operation_order=(form == "reference-order"))
    golden = case.golden
    assert res.Npv == pytest.approx(golden["npv"], rel=REL)
    assert round(abs(res.Npv - golden["npv"]), 6) == 0  # the reference's own assertion (assertAlmostEqual, places=6)
    assert res.ValuationSimStandardError == pytest.approx(golden["std_error"], rel=REL)
    assert notified[-1] == 1.0 and len(notified) == 245  # test_multi_factor.py:100-101
    periods = case.flat.periods
    deltas = np.array([res.Deltas[p] for p in periods])
    profile = np.array([[sp.Inventory, sp.InjectWithdrawVolume, sp.CmdtyConsumed, sp.InventoryLoss, sp.PeriodPv]
                        for sp in (res.ExpectedStorageProfile[p] for p in periods)])
    trigger = []
    for p in periods[:-1]:
        t = res.TriggerPrices[p]
        trigger.append((t.HasInjectPrice, t.MaxInjectVolume, t.MaxInjectTriggerPrice, t.HasWithdrawPrice, t.MaxWithdrawVolume,
                        t.MaxWithdrawTriggerPrice))
    worst = reference_goldens.compare_with_csvs(name, deltas, profile, trigger, rel=REL)
    assert res.InventoryBySim.shape == (123, 500) and res.RegressionSpotPriceSim.shape == (123, 500)  # :104-110
    print(f"{name} [{form}]: npv {res.Npv!r} vs golden {golden['npv']!r} (rel {abs(res.Npv - golden['npv']) / golden['npv']:.2e}), "
          f"std error {res.ValuationSimStandardError!r} vs {golden['std_error']!r}; worst relative differences vs the reference's CSVs {worst}")
