"""Measurement-set readers for the invert hot path.

The hot path only needs the duck-typed reader protocol of SURVEY.md appendix B
(reference: src/ska_sdp_cip/measurement_set.py:130-358): ``visibilities()``, ``flags()``,
``weights()``, ``uvw()``, ``channel_frequencies()``, ``partition()`` and the bound properties.
Chunked reads stay in host Python (north star); two implementations share the bound logic:

* ``ArrayMeasurementSetReader`` -- arrays already in host memory (synthetic data, tests, bench);
* ``MeasurementSetReader``      -- casacore-backed, same constructor as the reference; casacore is
  imported lazily because it is optional (absent in the build container).
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import Iterator, Optional, Union

import numpy as np
from numpy.typing import NDArray


def balanced_chunk_sizes(n: int, k: int) -> Iterator[int]:
    """Sizes of ``k`` chunks of ``n`` items, as even as possible: the first ``n % k`` chunks get one
    extra item (reference: measurement_set.py:361-376, golden-tested bounds)."""
    if n <= 0:
        raise ValueError("n must be > 0")
    if k <= 0 or k > n:
        raise ValueError("k must be > 0 and <= n")
    base, extra = divmod(n, k)
    for i in range(k):
        yield base + (1 if i < extra else 0)


def balanced_chunk_bounds(start: int, end: int, k: int) -> Iterator[tuple[int, int]]:
    """(start, end) index pairs of the balanced chunks of ``[start, end)`` (measurement_set.py:379-391)."""
    lo = start
    for size in balanced_chunk_sizes(end - start, k):
        yield lo, lo + size
        lo += size


class _BoundedReader:
    """Row / channel reading bounds and partitioning shared by all readers
    (reference: measurement_set.py:168-277)."""

    _total_rows: int
    _total_channels: int

    def _init_bounds(self, total_rows: int, total_channels: int) -> None:
        self._total_rows = int(total_rows)
        self._total_channels = int(total_channels)
        self._row_start, self._row_end = 0, self._total_rows
        self._channel_start, self._channel_end = 0, self._total_channels

    @property
    def row_start(self) -> int:
        return self._row_start

    @property
    def row_end(self) -> int:
        return self._row_end

    @property
    def num_data_rows(self) -> int:
        return self._row_end - self._row_start

    @property
    def channel_start(self) -> int:
        return self._channel_start

    @property
    def channel_end(self) -> int:
        return self._channel_end

    @property
    def num_channels(self) -> int:
        return self._channel_end - self._channel_start

    def set_row_bounds(self, row_start: int, row_end: int) -> None:
        self._row_start = max(row_start, 0)
        self._row_end = min(row_end, self._total_rows)

    def set_channel_bounds(self, channel_start: int, channel_end: int) -> None:
        self._channel_start = max(channel_start, 0)
        self._channel_end = min(channel_end, self._total_channels)

    def _clone(self):
        raise NotImplementedError

    def partition(self, row_chunks: int, freq_chunks: int) -> list:
        """Row-major, frequency-minor list of readers covering balanced chunks."""
        if not 1 <= row_chunks <= self.num_data_rows:
            raise ValueError("Number of row chunks must be within [1, total data rows]")
        if not 1 <= freq_chunks <= self.num_channels:
            raise ValueError("Number of row chunks must be within [1, total freq channels]")
        out = []
        for r0, r1 in balanced_chunk_bounds(self.row_start, self.row_end, row_chunks):
            for c0, c1 in balanced_chunk_bounds(self.channel_start, self.channel_end, freq_chunks):
                chunk = self._clone()
                chunk.set_row_bounds(r0, r1)
                chunk.set_channel_bounds(c0, c1)
                out.append(chunk)
        return out


class UnsupportedMeasurementSetLayout(Exception):
    """Raised when a MeasurementSet deviates from the supported layout (one spectral window, one field,
    one 4-correlation polarisation setup) -- same name as the reference's (measurement_set.py:12-16)."""


class ArrayMeasurementSetReader(_BoundedReader):
    """Reader over arrays held in host memory, with the reference reader's interface.

    vis complex64 (nrow, nchan, 4), flags bool (nrow, nchan, 4), weights float32 (nrow, nchan, 4),
    uvw float64 (nrow, 3), freqs float64 (nchan,).
    """

    def __init__(self, uvw: NDArray, freqs: NDArray, vis: NDArray, flags: NDArray, weights: NDArray,
                 path: Union[str, os.PathLike] = "in-memory.ms") -> None:
        nrow, nchan = vis.shape[:2]
        if vis.shape != (nrow, nchan, 4) or flags.shape != vis.shape or weights.shape != vis.shape:
            raise ValueError("vis, flags and weights must all have shape (nrow, nchan, 4)")
        if uvw.shape != (nrow, 3) or freqs.shape != (nchan,):
            raise ValueError("uvw must be (nrow, 3) and freqs (nchan,)")
        self._uvw, self._freqs, self._vis, self._flags, self._weights = uvw, freqs, vis, flags, weights
        self._path = Path(path).resolve()
        self._init_bounds(nrow, nchan)

    @property
    def path(self) -> Path:
        return self._path

    def _clone(self) -> "ArrayMeasurementSetReader":
        return ArrayMeasurementSetReader(self._uvw, self._freqs, self._vis, self._flags, self._weights, self._path)

    def _rows(self) -> slice:
        return slice(self.row_start, self.row_end)

    def _chans(self) -> slice:
        return slice(self.channel_start, self.channel_end)

    def channel_frequencies(self) -> NDArray:
        return self._freqs[self._chans()]

    def uvw(self) -> NDArray:
        return self._uvw[self._rows()]

    def flags(self) -> NDArray:
        return self._flags[self._rows(), self._chans()]

    def visibilities(self) -> NDArray:
        return self._vis[self._rows(), self._chans()]

    def weights(self) -> NDArray:
        return self._weights[self._rows(), self._chans()]


class MeasurementSetReader(_BoundedReader):
    """casacore-backed reader of a MeasurementSet v2 with the reference's constructor and methods
    (measurement_set.py:130-358).  Single spectral window, 4 correlations assumed, as there."""

    def __init__(self, path: Union[str, os.PathLike], *, validate_layout: bool = True) -> None:
        self._path = Path(path).resolve()
        if not self._path.is_dir():
            raise FileNotFoundError(f"Measurement set {str(self._path)!r} does not exist")
        with self._table("MAIN") as tbl:
            nrows = tbl.nrows()
        with self._table("SPECTRAL_WINDOW") as tbl:
            if validate_layout and tbl.nrows() != 1:
                raise UnsupportedMeasurementSetLayout("MeasurementSet must have exactly one spectral window")
            nchan = len(tbl.getcol("CHAN_FREQ")[0])
        if validate_layout:
            with self._table("POLARIZATION") as tbl:
                corr = list(tbl.getcol("CORR_TYPE")[0])
                if tbl.nrows() != 1 or corr not in ([9, 10, 11, 12], [5, 6, 7, 8]):
                    raise UnsupportedMeasurementSetLayout(
                        "MeasurementSet must hold one linear or circular 4-correlation setup")
            with self._table("FIELD") as tbl:
                if tbl.nrows() != 1:
                    raise UnsupportedMeasurementSetLayout("MeasurementSet must have exactly one field")
        self._init_bounds(nrows, nchan)

    def _table(self, name: str):
        try:
            from casacore.tables import table  # pylint:disable=import-outside-toplevel
        except ImportError as err:  # pragma: no cover - casacore is optional
            raise ImportError("python-casacore is required to read MeasurementSets from disk; "
                              "use ArrayMeasurementSetReader for data in memory") from err
        p = str(self._path) if name == "MAIN" else f"{self._path}::{name}"
        return table(p, ack=False, readonly=True)

    @property
    def path(self) -> Path:
        return self._path

    def _clone(self) -> "MeasurementSetReader":
        return MeasurementSetReader(self._path, validate_layout=False)

    def channel_frequencies(self) -> NDArray:
        with self._table("SPECTRAL_WINDOW") as tbl:
            return tbl.getcolslice("CHAN_FREQ", blc=self.channel_start, trc=self.channel_end - 1)[0]

    def uvw(self) -> NDArray:
        with self._table("MAIN") as tbl:
            return tbl.getcol("UVW", startrow=self.row_start, nrow=self.num_data_rows)

    def _cube(self, column: str) -> NDArray:
        with self._table("MAIN") as tbl:
            return tbl.getcolslice(column, blc=(self.channel_start, 0), trc=(self.channel_end - 1, 3),
                                   startrow=self.row_start, nrow=self.num_data_rows)

    def flags(self) -> NDArray:
        return self._cube("FLAG")

    def visibilities(self) -> NDArray:
        return self._cube("DATA")

    def weights(self) -> NDArray:
        try:
            return self._cube("WEIGHT_SPECTRUM")
        except RuntimeError:  # no WEIGHT_SPECTRUM column: repeat WEIGHT along frequency
            with self._table("MAIN") as tbl:
                data = tbl.getcolslice("WEIGHT", blc=0, trc=3, startrow=self.row_start, nrow=self.num_data_rows)
            nrow, npol = data.shape
            return data.reshape(nrow, 1, npol).repeat(self.num_channels, axis=1)
