"""CPU restatement of the reference's spot-price simulation chain (TEST INFRASTRUCTURE — see oracle/oracle.py's header).

The reference simulates with un-vendored third-party code: Cmdty.Core 0.2.0's
MultiFactorSpotPriceSimulator<T> driven by its MersenneTwisterGenerator, which wraps MathNet.Numerics' MersenneTwister and
Normal.Samples (call sites: LsmcValuation/LsmcValuationParameters.cs:164-196).  Neither library is in /root/reference, so
their published algorithms are restated here and PINNED on the reference's own golden vectors:

  * src/Cmdty.Storage.Python/tests/test_multi_factor_spot_sim.py:64-89 — twelve 17-digit spot prices (3 factors, seed 12):
    reproduced to the last digit (tests/test_reference_pinning.py);
  * src/Cmdty.Storage.Python/tests/test_multi_factor.py:99-124 — the LSMC goldens that these paths feed.

The chain, as pinned:
  1. MT19937 seeded with init_genrand(seed) (Matsumoto & Nishimura reference code, as MathNet's MersenneTwister(int seed)).
  2. MathNet's NextDouble(): genrand_int32() * 2^-32.
  3. MathNet's Normal.Samples(rnd, double[] values, 0, 1): draw n = ceil(len * 4/pi) uniforms (made even) in ONE block, turn
     consecutive pairs into normals by the polar (Marsaglia) transform, both outputs used in order, rejected pairs skipped; if
     the block runs out, continue pair by pair.  What is left of a block is discarded.
  4. Cmdty.Core: one such array per simulation, laid out [period][factor]; with `antithetic` every second simulation re-uses
     the previous simulation's array with the sign flipped.
  5. Markov factors by the exact Ornstein-Uhlenbeck transition: x_i <- x_i exp(-a_i dt) + (L z)_i, L the lower Cholesky factor
     of the increment covariance rho_ij (1 - exp(-(a_i + a_j) dt)) / (a_i + a_j); spot S = F(0,T) exp(sum_i vol_i(T) x_i - Var/2).
"""
import math

import numpy as np


class MathNetMersenneTwister:
    def __init__(self, seed):
        self._bits = np.random.MT19937()
        self._bits._legacy_seeding(int(seed))  # init_genrand(seed)

    def next_double(self):
        return float(self._bits.random_raw(1)[0]) / 4294967296.0

    def next_doubles(self, n):
        return self._bits.random_raw(n).astype(np.float64) / 4294967296.0


def _polar(a, b):
    v1 = 2.0 * a - 1.0
    v2 = 2.0 * b - 1.0
    r = v1 * v1 + v2 * v2
    if r >= 1.0 or r == 0.0:
        return None
    fac = math.sqrt(-2.0 * math.log(r) / r)
    return v1 * fac, v2 * fac


def normal_samples(rnd, count):
    """MathNet.Numerics.Distributions.Normal.Samples(rnd, new double[count], 0.0, 1.0)."""
    values = []
    if count == 0:
        return values
    n = int(math.ceil(count * 4 * (1.0 / math.pi)))
    if n % 2:
        n += 1
    u = rnd.next_doubles(n)
    i = 0
    while i < n and len(values) < count:
        pair = _polar(u[i], u[i + 1])
        i += 2
        if pair is None:
            continue
        values.append(pair[0])
        if len(values) < count:
            values.append(pair[1])
    while len(values) < count:
        pair = _polar(rnd.next_double(), rnd.next_double())
        if pair is None:
            continue
        values.append(pair[0])
        if len(values) < count:
            values.append(pair[1])
    return values


def _integral(a, t):
    return t if a == 0.0 else (1.0 - math.exp(-a * t)) / a


def simulate(mean_reversion, vol_by_period, corr, times, forward, num_sims, seed, antithetic):
    """Returns spot [rows][N] and factors [F][rows][N].  vol_by_period: [F][rows]; times: Act/365 year fractions from the
    current date to every simulated period; forward: [rows]."""
    mr = np.asarray(mean_reversion, dtype=np.float64)
    vol = np.asarray(vol_by_period, dtype=np.float64)
    corr = np.asarray(corr, dtype=np.float64)
    F, rows = vol.shape
    chol, decay, half_var = [], [], []
    t_prev = 0.0
    for k in range(rows):
        t = times[k]
        dt = t - t_prev
        cov = np.array([[corr[i, j] * _integral(mr[i] + mr[j], dt) for j in range(F)] for i in range(F)])
        chol.append(np.linalg.cholesky(cov))
        decay.append(np.exp(-mr * dt))
        half_var.append(0.5 * sum(vol[i, k] * vol[j, k] * corr[i, j] * _integral(mr[i] + mr[j], t) for i in range(F) for j in range(F)))
        t_prev = t
    rnd = MathNetMersenneTwister(seed)
    spot = np.zeros((rows, num_sims))
    factors = np.zeros((F, rows, num_sims))
    z = None
    for s in range(num_sims):
        if antithetic and s % 2 == 1:
            z = -z
        else:
            z = np.array(normal_samples(rnd, rows * F)).reshape(rows, F)
        x = np.zeros(F)
        for k in range(rows):
            x = x * decay[k] + chol[k] @ z[k]
            factors[:, k, s] = x
            spot[k, s] = forward[k] * math.exp(float(vol[:, k] @ x) - half_var[k])
    return spot, factors
