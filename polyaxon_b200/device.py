"""Process-wide device context: one C-ABI handle per (process, GPU).

The caller of the suggestion path is a Celery prefork worker (``cmd/celery_worker``): one thread
per process, several processes per host.  Each process lazily opens one handle on the GPU named by
``PLX_BO_DEVICE`` (default: ``LOCAL_RANK`` under torchrun, else 0) and keeps it for its lifetime.
"""
import os

from polyaxon_b200 import _lib

_handles = {}


def default_device():
    return int(os.environ.get('PLX_BO_DEVICE', os.environ.get('LOCAL_RANK', '0')))


def get_handle(device=None):
    device = default_device() if device is None else int(device)
    h = _handles.get(device)
    if h is None:
        h = _handles[device] = _lib.Handle(device)
    return h


def close_all():
    for h in _handles.values():
        h.close()
    _handles.clear()
