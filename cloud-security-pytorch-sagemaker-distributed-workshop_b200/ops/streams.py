"""Auxiliary CUDA stream for work that is off the backward critical path.

In backward, a layer's weight gradient (wgrad) is needed only by the bucket all-reduce and the
optimizer, while its input gradient (dgrad) gates the next layer.  The reference's resnet18 runs at
8x8..1x1 spatial, so a single kernel fills a fraction of the 148 SMs: wgrad kernels are forked onto this
stream and run concurrently with the dgrad chain (in a captured CUDA graph they become parallel
branches).  ``join`` makes the current stream wait for everything forked so far.
"""
from __future__ import annotations

import contextlib
import os

import torch

_AUX = {}
_DIRTY = {}
_MAIN = {}
ENABLED = [os.environ.get("B200DDP_AUX_STREAM", "1") not in ("0", "false", "off")]


def _key(device):
    return device.index if device.index is not None else torch.cuda.current_device()


def aux_stream(device) -> torch.cuda.Stream:
    k = _key(device)
    s = _AUX.get(k)
    if s is None:
        s = torch.cuda.Stream(device=device)
        _AUX[k] = s
    return s


@contextlib.contextmanager
def forked(device, *tensors):
    """Run the body on the aux stream, ordered after everything queued on the current stream.
    ``tensors`` are the inputs the body reads (kept alive for the aux stream's use)."""
    if not ENABLED[0]:
        yield
        return
    cur = torch.cuda.current_stream(device)
    aux = aux_stream(device)
    ev = torch.cuda.Event()
    ev.record(cur)
    aux.wait_event(ev)
    for t in tensors:
        if t is not None:
            t.record_stream(aux)
    _DIRTY[_key(device)] = True
    _MAIN[_key(device)] = cur
    with torch.cuda.stream(aux):
        yield


def join(device) -> None:
    """Current stream waits for all forked work."""
    k = _key(device)
    if _DIRTY.get(k):
        ev = torch.cuda.Event()
        ev.record(aux_stream(device))
        torch.cuda.current_stream(device).wait_event(ev)
        _DIRTY[k] = False


def producers(device):
    """Streams that may hold not-yet-finished gradient producers right now: the stream the step runs on
    and the aux stream (whichever of them is not the current stream must be waited on explicitly)."""
    k = _key(device)
    if not _DIRTY.get(k):
        return []
    cur = torch.cuda.current_stream(device)
    return [s for s in (_MAIN.get(k), _AUX.get(k)) if s is not None and s != cur]
