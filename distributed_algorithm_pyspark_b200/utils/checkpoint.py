"""Checkpoint / resume.

The reference has no checkpointing at all (``.cache()`` is memoisation, results are
only printed).  Long runs on a GPU box want one, so every iterative driver here can write its full
optimiser state every ``k`` iterations and continue from it.  Because all sampling is a pure function of
``(seed, iteration, row)`` (Spark's ``sample(False, f, 42 + t)``, ssgd.py:97, and the Philox device
sampler), a resumed run follows the uninterrupted trajectory bit for bit.

Files are ``.npz`` archives written atomically (temp file + ``os.replace``): a crash while writing never
leaves a truncated checkpoint behind.  State that is replicated across ranks is written by rank 0;
rank-local state (e.g. EASGD replicas of the benchmark engine) goes to ``<path>.rank<r>``.
"""
from __future__ import annotations

import json
import os
import tempfile
from typing import Any, Dict, Optional, Tuple

import numpy as np

FORMAT_VERSION = 1


def _to_numpy(v):
    try:
        import torch
        if isinstance(v, torch.Tensor):
            v = v.detach()
            if v.dtype == torch.bfloat16:
                v = v.float()
            return v.cpu().numpy()
    except ImportError:                                   # pragma: no cover
        pass
    return np.asarray(v)


def save_checkpoint(path: str, state: Dict[str, Any], meta: Optional[Dict[str, Any]] = None) -> str:
    """Atomically write ``state`` (arrays / tensors / scalars; ``None`` entries are skipped) and JSON ``meta``."""
    path = os.fspath(path)
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    arrays = {k: _to_numpy(v) for k, v in state.items() if v is not None}
    header = dict(meta or {})
    header["format_version"] = FORMAT_VERSION
    arrays["__meta__"] = np.frombuffer(json.dumps(header, sort_keys=True).encode(), dtype=np.uint8)
    fd, tmp = tempfile.mkstemp(prefix=".ckpt_", suffix=".tmp", dir=d)
    try:
        with os.fdopen(fd, "wb") as f:
            np.savez(f, **arrays)
            f.flush()
            os.fsync(f.fileno())
        os.replace(tmp, path)
    except BaseException:
        if os.path.exists(tmp):
            os.unlink(tmp)
        raise
    return path


def load_checkpoint(path: str) -> Tuple[Dict[str, np.ndarray], Dict[str, Any]]:
    """Returns ``(state, meta)``; raises ``FileNotFoundError`` if absent, ``ValueError`` on a foreign file."""
    with np.load(os.fspath(path), allow_pickle=False) as z:
        if "__meta__" not in z.files:
            raise ValueError(f"{path}: not a checkpoint written by this package")
        meta = json.loads(bytes(z["__meta__"]).decode())
        if meta.get("format_version") != FORMAT_VERSION:
            raise ValueError(f"{path}: checkpoint format {meta.get('format_version')} != {FORMAT_VERSION}")
        state = {k: z[k] for k in z.files if k != "__meta__"}
    return state, meta


def rank_path(path: str, rank: int) -> str:
    return f"{os.fspath(path)}.rank{rank}"


def check_meta(meta: Dict[str, Any], **expected) -> None:
    """A checkpoint may only be resumed by the same algorithm / problem shape."""
    for k, v in expected.items():
        if k in meta and meta[k] != v:
            raise ValueError(f"checkpoint mismatch: {k} = {meta[k]!r} in the file, {v!r} requested")
