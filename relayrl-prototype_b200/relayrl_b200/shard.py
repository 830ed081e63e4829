"""Episode-aligned sharding for the data-parallel learner.

Episodes are the independent units of the scan, so ranks own contiguous runs of WHOLE episodes
(no cross-GPU carry) balanced by cumulative transitions.  Pure host logic (numpy)."""
import numpy as np


def shard_episodes(offsets, nranks):
    """offsets: int64 [E+1] successive path_start_idx values (offsets[0] == 0).
    Returns episode boundaries `eb` (nranks+1 ints): rank r owns episodes [eb[r], eb[r+1]).
    Cut points are the episode ends closest to the ideal transition split r*N/nranks."""
    offsets = np.asarray(offsets, np.int64)
    E, N = len(offsets) - 1, int(offsets[-1])
    eb = [0]
    for r in range(1, nranks):
        target = N * r / nranks
        k = int(np.searchsorted(offsets, target, side="left"))
        if k > 0 and (k > E or abs(offsets[k - 1] - target) <= abs(offsets[min(k, E)] - target)):
            k -= 1
        eb.append(min(max(k, eb[-1]), E))
    eb.append(E)
    return np.asarray(eb, np.int64)


def shard_rows(offsets, eb, rank):
    """Transition range [lo, hi) and the rank-local offsets of rank `rank`."""
    offsets = np.asarray(offsets, np.int64)
    lo, hi = int(offsets[eb[rank]]), int(offsets[eb[rank + 1]])
    return lo, hi, offsets[eb[rank]:eb[rank + 1] + 1] - lo


def merge_offsets(local_offsets_per_rank):
    """Concatenates rank-local offsets back into the global array (parity check: the result must
    equal the single-GPU offsets bit for bit)."""
    out, base = [np.zeros(1, np.int64)], 0
    for lo in local_offsets_per_rank:
        lo = np.asarray(lo, np.int64)
        out.append(lo[1:] + base)
        base += int(lo[-1])
    return np.concatenate(out)
