"""Process-wide device context: one C-ABI handle per (process, GPU, worker slot).

The caller of the suggestion path is a Celery prefork worker (``cmd/celery_worker``): one thread
per process, several processes per host.  Each process lazily opens one handle on the GPU named by
``PLX_BO_DEVICE`` (default: ``LOCAL_RANK`` under torchrun, else 0) and keeps it for its lifetime.

``polyaxon_b200.groups.suggest_groups`` runs several independent groups at once from worker threads;
each of those threads names a slot (``use_slot``) and gets a handle of its own -- own stream, own
workspace, own resident candidate matrix -- so their kernels overlap on the GPU.  A handle is never
shared between threads (the C ABI is not thread-safe per handle).
"""
import os
import threading

from polyaxon_b200 import _lib

_handles = {}
_lock = threading.Lock()
_tls = threading.local()


def default_device():
    return int(os.environ.get('PLX_BO_DEVICE', os.environ.get('LOCAL_RANK', '0')))


def use_slot(slot):
    """Name the handle slot of the calling thread (0 = the process-wide default)."""
    _tls.slot = int(slot)


def get_handle(device=None):
    device = default_device() if device is None else int(device)
    key = (device, getattr(_tls, 'slot', 0))
    with _lock:
        h = _handles.get(key)
        if h is None:
            h = _handles[key] = _lib.Handle(device)
            # PLX_BO_RESERVE="n,d,runs,seeds,candidates": allocate the workspaces of that shape when the worker's
            # handle is opened, so that no call of a group pays for device allocations (plx_bo_reserve)
            # PLX_BO_BLOCKING_SYNC=1 (set by groups.suggest_groups when its worker threads outnumber the cores the
            # process may use): wait for the GPU asleep on an event instead of spinning
            if os.environ.get('PLX_BO_BLOCKING_SYNC', '0') == '1':
                h.set_option('blocking_sync', 1)
            spec = os.environ.get('PLX_BO_RESERVE')
            if spec:
                h.reserve(*[int(v) for v in spec.split(',')])
    return h


def open_handles():
    with _lock:
        return list(_handles.values())


def close_all():
    with _lock:
        for h in _handles.values():
            h.close()
        _handles.clear()
