"""ORACLE (test infrastructure): balanced chunk bounds (P7).

Restates balanced_chunk_sizes / balanced_chunk_bounds (src/ska_sdp_cip/measurement_set.py:361-391)
and the row-major / frequency-minor order of MeasurementSetReader.partition (:234-277).
PINNED by the reference's golden values (tests/test_measurement_set_partition_indices.py:33-62:
74,214 rows x 4 channels for (1,1), (2,3), (5,1)), copied as numbers into tests/test_host_logic.py.
"""
from __future__ import annotations


def chunk_bounds(start: int, end: int, k: int) -> list[tuple[int, int]]:
    n = end - start
    if n <= 0:
        raise ValueError("n must be > 0")
    if k <= 0 or k > n:
        raise ValueError("k must be > 0 and <= n")
    q, r = divmod(n, k)
    out, lo = [], start
    for i in range(k):
        hi = lo + q + (1 if i < r else 0)
        out.append((lo, hi))
        lo = hi
    return out


def partition_bounds(nrow: int, nchan: int, row_chunks: int, freq_chunks: int):
    """[(row_start, row_end, chan_start, chan_end)] in the reference's order."""
    return [(r0, r1, c0, c1) for r0, r1 in chunk_bounds(0, nrow, row_chunks)
            for c0, c1 in chunk_bounds(0, nchan, freq_chunks)]
