"""``ska-sdp-cip``: invert a MeasurementSet into a dirty image saved as ``.npy`` -- same command line as the
reference app (src/ska_sdp_cip/apps/pipeline_app.py:16-120).  Without ``--dask-scheduler`` the invert runs in
this process on the current GPU (under ``torchrun`` every rank takes its share of the chunks and the images
are combined with one NCCL reduction); with a scheduler address the chunks fan out as dask tasks and the task
stream is saved to ``task-list.json`` like the reference does."""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import numpy as np

from .. import MeasurementSetReader, __version__, dask_invert_measurement_set, invert_measurement_set
from ..task_metrics import TaskMetrics


def get_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Launch the SKA continuum imaging pipeline",
                                     formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("--version", action="version", version=__version__)
    parser.add_argument("measurement_set", type=Path, help="Path to MeasurementSet v2")
    parser.add_argument("output_image", type=Path, help="Path to output image, which is saved as a numpy array")
    imaging = parser.add_argument_group("imaging")
    imaging.add_argument("-n", "--num-pixels", type=int, required=True, help="Number of pixels across the image")
    imaging.add_argument("-p", "--pixel-size", type=float, required=True,
                         help="Pixel size in arcseconds at the image centre")
    dist = parser.add_argument_group("dask distribution")
    dist.add_argument("-d", "--dask-scheduler", type=str, default=None,
                      help="Optional address of a dask scheduler to use for distribution")
    dist.add_argument("-rc", "--row-chunks", type=int, default=1, help="Number of row chunks to use for distribution")
    dist.add_argument("-fc", "--freq-chunks", type=int, default=None,
                      help="Number of frequency chunks to use for distribution. "
                           "If None, set this to the number of dask workers.")
    return parser


def run_program(cli_args: list[str], *, reader_factory=MeasurementSetReader) -> None:
    """Run the app (the function the reference's tests call).  ``reader_factory`` lets tests substitute an
    in-memory reader for the casacore-backed one."""
    args = get_parser().parse_args(cli_args)
    mset = reader_factory(args.measurement_set)
    if args.dask_scheduler is None:
        img = invert_measurement_set(mset, num_pixels=args.num_pixels, pixel_size_asec=args.pixel_size)
    else:
        from dask.distributed import Client, get_task_stream, performance_report  # pylint:disable=import-outside-toplevel
        client = Client(args.dask_scheduler)
        with get_task_stream(client) as stream, performance_report(filename="dask-report.html"):
            img = dask_invert_measurement_set(mset, client, num_pixels=args.num_pixels,
                                              pixel_size_asec=args.pixel_size, row_chunks=args.row_chunks,
                                              freq_chunks=args.freq_chunks)
        TaskMetrics(stream.data).save_json("task-list.json", indent=4, sort_keys=True)
    np.save(args.output_image.with_suffix(".npy"), img)


def main() -> None:
    run_program(sys.argv[1:])


if __name__ == "__main__":
    sys.exit(main())
