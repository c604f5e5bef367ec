"""plx_acq_sweep_noalloc through the C ABI, against the ordinary entry point.

This file is collected last on purpose (a run that stops at the first failure has then already
run everything else).  The entry point is meant for hosts that may not allocate device memory
inside a library call -- a process that has given the whole GPU to its own memory pool, or one
under a fixed MPS / container memory limit -- so the test also checks that free device memory does
not move across a call.  Two calls in a row share the static scratch array; the second one runs on
different candidates and a different observation count so that stale contents would show.

Tolerance: arg-max index exact, best value 1e-9 relative (the tolerances of test_gpu_parity.py).
"""
import numpy as np
import pytest

from oracle import gp_numpy as G
from oracle.reference_path import synthetic_problem

pytestmark = pytest.mark.gpu


def _run_pair(handle, torch, n, d, n_cand, seed, acq, kappa, eps):
    from polyaxon_b200 import _lib
    X, y, _ = synthetic_problem(n, d, seed=seed)
    theta = 0.4
    handle.set_data(X, y, _lib.KERNEL_MATERN, 2.5)
    handle.set_theta(theta)
    _, alpha, linv = handle.get_state(want_L=False, want_Linv=True)
    xc = np.random.RandomState(seed).uniform(0, 1, (n_cand, d))
    y_max = float(y.max())

    dev = torch.device('cuda', 0)
    xc_d = torch.from_numpy(xc).to(dev)
    x_d = torch.from_numpy(X).to(dev)
    a_d = torch.from_numpy(alpha).to(dev)
    z_d = torch.from_numpy(linv).to(dev)
    out_v = torch.zeros(1, dtype=torch.float64, device=dev)
    out_i = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()

    want_v, want_i, _ = handle.sweep_dev(xc_d.data_ptr(), n_cand, acq, y_max, kappa=kappa, eps=eps, idx_base=5)

    stream = torch.cuda.current_stream().cuda_stream
    free_before = torch.cuda.mem_get_info()[0]
    _lib.sweep_noalloc(0, stream, xc_d.data_ptr(), n_cand, x_d.data_ptr(), a_d.data_ptr(), z_d.data_ptr(),
                       n, n, d, _lib.KERNEL_MATERN, 2.5, theta, acq, y_max, out_v.data_ptr(), out_i.data_ptr(),
                       kappa=kappa, eps=eps, idx_base=5)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return (out_v.item(), out_i.item()), (want_v, want_i), free_before - free_after


def test_noalloc_sweep_matches_the_handle_entry_point_twice():
    import torch
    from polyaxon_b200 import _lib
    handle = _lib.Handle(0)
    try:
        # first call on this device also performs the one-time set-up
        got, want, _ = _run_pair(handle, torch, n=200, d=6, n_cand=9000, seed=11, acq=_lib.ACQ_EI, kappa=0.0, eps=0.01)
        assert got[1] == want[1]
        assert got[0] == pytest.approx(want[0], rel=1e-9)
        # second call: fewer observations (other padding), other candidates, other acquisition
        got, want, moved = _run_pair(handle, torch, n=37, d=3, n_cand=700, seed=12, acq=_lib.ACQ_UCB, kappa=2.0, eps=0.0)
        assert got[1] == want[1]
        assert got[0] == pytest.approx(want[0], rel=1e-9)
        assert moved == 0, 'device memory changed across a no-alloc call'
    finally:
        handle.close()


def test_noalloc_sweep_refuses_sizes_beyond_the_static_scratch():
    import torch
    from polyaxon_b200 import _lib
    dev = torch.device('cuda', 0)
    buf = torch.zeros(8, dtype=torch.float64, device=dev)
    idx = torch.zeros(1, dtype=torch.int64, device=dev)
    with pytest.raises(_lib.PlxError, match='max 256 observations'):
        _lib.sweep_noalloc(0, None, buf.data_ptr(), 1, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), 300, 300, 2,
                           _lib.KERNEL_RBF, 1.5, 0.0, _lib.ACQ_UCB, 0.0, buf.data_ptr(), idx.data_ptr())
