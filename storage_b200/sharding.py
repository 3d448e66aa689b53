"""Host-side helpers for path sharding: which slice of the paths a rank owns and how per-rank statistics
combine.  The library does the same combination on the device (all-gather, then sums in rank order); these helpers are
what a multi-process host (bench.py under torchrun) uses around the C ABI, and what the CPU gloo tests exercise."""
import math


def shard_bounds(total_paths, world, rank):
    """Contiguous slice [begin, end) of rank `rank`; slices differ by at most one path and start on an even path so that
    antithetic pairs (2j, 2j+1) are never split across ranks."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    pairs = (total_paths + 1) // 2
    base, extra = divmod(pairs, world)
    begin_pair = rank * base + min(rank, extra)
    end_pair = begin_pair + base + (1 if rank < extra else 0)
    return min(2 * begin_pair, total_paths), min(2 * end_pair, total_paths)


def merge_npv(rank_sums, total_paths):
    """NPV from per-rank sums of PvBySim, summed in rank order (LsmcStorageValuation.cs:617)."""
    total = 0.0
    for s in rank_sums:
        total += s
    return total / total_paths


def merge_standard_error(rank_sq_dev_sums, total_paths):
    """Sample standard error (N-1) from per-rank sums of (pv - npv)^2 around the GLOBAL mean (LsmcStorageValuation.cs:618)."""
    total = 0.0
    for s in rank_sq_dev_sums:
        total += s
    variance = total / (total_paths - 1) if total_paths > 1 else 0.0
    return math.sqrt(variance) / math.sqrt(total_paths)


def merge_regression_sums(rank_sums):
    """Element-wise sum of per-rank regression sums (Q^T V partials) in rank order: identical on every rank."""
    out = [0.0] * len(rank_sums[0])
    for sums in rank_sums:
        for i, v in enumerate(sums):
            out[i] += v
    return out
