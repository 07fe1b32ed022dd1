"""B200-native drop-in for the invert hot path of ``ska_sdp_cip``
(reference public surface: src/ska_sdp_cip/__init__.py:1-10)."""
from ._version import __version__
from .invert import dask_invert_measurement_set, invert_measurement_set
from .measurement_set import MeasurementSetReader

__all__ = [
    "__version__",
    "MeasurementSetReader",
    "dask_invert_measurement_set",
    "invert_measurement_set",
]
