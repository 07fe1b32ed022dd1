"""Drop-in for ``ska_sdp_cip.task_metrics`` (reference: src/ska_sdp_cip/task_metrics.py:10-135): the per-task
records of a dask task stream (``get_task_stream().data``) as a sequence that exports to JSON.  The same
container also takes the stage reports of the GPU library (``TaskMetrics.from_stage_reports``), so that a run
without dask still leaves a ``task-list.json`` with one entry per library call."""
from __future__ import annotations

import collections.abc
import json
import os
from dataclasses import dataclass, field
from typing import Iterable, Union

_EXPORTED = ("key", "worker", "status", "start", "stop", "name", "duration")


@dataclass
class Task:
    """One finished task as reported by the dask scheduler (task_metrics.py:10-86): ``key`` unique id,
    ``worker`` address, ``status``, ``start`` / ``stop`` UNIX timestamps; derived: ``name`` = key up to its
    last "-", ``duration`` = stop - start in seconds."""

    key: str
    worker: str
    status: str
    start: float
    stop: float
    name: str = field(init=False)
    duration: float = field(init=False)

    def __post_init__(self) -> None:
        self.name = self.key.rsplit("-", maxsplit=1)[0]
        self.duration = self.stop - self.start

    def as_dict(self) -> dict:
        return {k: getattr(self, k) for k in _EXPORTED}

    @classmethod
    def _from_task_stream_entry(cls, entry: dict) -> "Task":
        """``entry["startstops"]`` lists the actions ("compute", "transfer", ...) of the task with their own
        start/stop stamps; the task spans from the earliest start to the latest stop."""
        spans = entry["startstops"]
        return cls(key=entry["key"], worker=entry["worker"], status=entry["status"],
                   start=min(s["start"] for s in spans), stop=max(s["stop"] for s in spans))


class TaskMetrics(collections.abc.Sequence):
    """Sequence of ``Task`` parsed from dask's task stream data; ``to_json`` / ``save_json`` give a list of
    dicts that ``pandas.read_json`` loads (task_metrics.py:88-135)."""

    def __init__(self, task_stream_data: Iterable[dict]) -> None:
        self._tasks = [Task._from_task_stream_entry(e) for e in task_stream_data]

    @classmethod
    def from_stage_reports(cls, reports: Iterable[dict], *, worker: str = "cuda:0", start: float = 0.0
                           ) -> "TaskMetrics":
        """One task per stage of each ``cip_report`` dict (keys ``ms_scan``, ``ms_bin``, ...), laid end to end
        from ``start``: the GPU-side counterpart of the dask task stream."""
        entries, t = [], float(start)
        for i, rep in enumerate(reports):
            for stage, ms in rep.items():
                if not stage.startswith("ms_") or stage == "ms_total" or not ms:
                    continue
                entries.append({"key": f"{stage[3:]}-{i}", "worker": worker, "status": "OK",
                                "startstops": ({"action": "compute", "start": t, "stop": t + ms * 1e-3},)})
                t += ms * 1e-3
        return cls(entries)

    def __len__(self) -> int:
        return len(self._tasks)

    def __getitem__(self, index):
        return self._tasks[index]

    def to_json(self, **kwargs) -> str:
        return json.dumps([t.as_dict() for t in self._tasks], **kwargs)

    def save_json(self, path: Union[str, os.PathLike], **kwargs) -> None:
        with open(path, "w", encoding="utf-8") as fh:
            fh.write(self.to_json(**kwargs))
