"""``ska-sdp-cip-reorder-uvw``: convert MSv2 visibilities to Stokes I and sort them into UVW tile chunk files --
the command line of the reference app (src/ska_sdp_cip/apps/uvw_reorder_app.py:12-113).  The scheduler
address is optional here: without it both passes of the reorder run in this process, binning and gathering
on the GPU."""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

from .. import MeasurementSetReader, __version__
from ..task_metrics import TaskMetrics
from ..uvw_tiling import reorder_by_uvw_tile


def get_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Convert MSv2 visibilities to Stokes I and sort them by UVW tile",
                                     formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("--version", action="version", version=__version__)
    parser.add_argument("measurement_set", type=Path, help="Path to MeasurementSet v2")
    parser.add_argument("-t", "--tile-size", nargs=3, type=float, required=True,
                        help="UVW tile size in units of wavelength, as a space-separated sequence of 3 "
                             "real-valued numbers")
    parser.add_argument("-d", "--dask-scheduler", type=str, default=None,
                        help="Address of a dask scheduler to use for distribution (default: run in-process)")
    parser.add_argument("-o", "--outdir", type=Path, default=Path.cwd(),
                        help="Output directory for the reordered data (and temporary files). Will be created if "
                             "it does not exist. Defaults to the current working directory.")
    parser.add_argument("-n", "--num-time-intervals", type=int, default=None,
                        help="Split the input data into this many time chunks between workers. "
                             "If None, a choice is made automatically.")
    parser.add_argument("-m", "--max-vis-per-chunk", type=int, default=5_000_000,
                        help="Maximum number of visibility samples to be stored per tile chunk.")
    return parser


def run_program(cli_args: list[str], *, reader_factory=MeasurementSetReader) -> list[Path]:
    args = get_parser().parse_args(cli_args)
    mset = reader_factory(args.measurement_set)
    outdir: Path = args.outdir
    outdir.mkdir(parents=True, exist_ok=True)
    kwargs = {"num_time_intervals": args.num_time_intervals, "max_vis_per_chunk": args.max_vis_per_chunk}
    if args.dask_scheduler is None:
        return reorder_by_uvw_tile(mset, tuple(args.tile_size), outdir, None, **kwargs)
    from dask.distributed import Client, get_task_stream, performance_report  # pylint:disable=import-outside-toplevel
    client = Client(args.dask_scheduler)
    with get_task_stream(client) as stream, performance_report(filename="dask-report.html"):
        paths = reorder_by_uvw_tile(mset, tuple(args.tile_size), outdir, client, **kwargs)
    TaskMetrics(stream.data).save_json("task-list.json", indent=4, sort_keys=True)
    return paths


def main() -> None:
    run_program(sys.argv[1:])


if __name__ == "__main__":
    sys.exit(main())
