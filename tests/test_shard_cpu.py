"""N>1 host logic on CPU (world_size-2 `gloo`): the partition rules, and the algebra the row-sharded fit
relies on — every rank streams only its rows, all-reduces S = X'Y once and [g | tt] once per component, and
replicates the K-sized step — reproduce the single-process oracle; fold-sharded CV assembles the same
residual block.  The numpy driver below restates pls_b200/csrc/fit.cu:fit_device for the test only."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import port
from pls_b200 import sharding
from tests.util import RTOL, relerr


def test_bounds_cover_exactly_once():
    for total in (0, 1, 7, 60, 1000, 1_000_000):
        for world in (1, 2, 3, 4, 8):
            b = sharding.bounds(total, world)
            assert b[0] == 0 and b[-1] == total and all(y >= x for x, y in zip(b, b[1:]))
            sizes = [y - x for x, y in zip(b, b[1:])]
            assert max(sizes) - min(sizes) <= 1
            assert sharding.row_shard(total, world - 1, world) == (b[-2], b[-1])
            assert sharding.fold_shard(total, 0, world) == (0, b[1])
    with pytest.raises(ValueError):
        sharding.bounds(5, 0)


def _allreduce(a):
    t = torch.from_numpy(np.ascontiguousarray(a))
    dist.all_reduce(t)
    return t.numpy()


def _sharded_fit(X, Y, A):
    """fit.cu's KERNEL_TYPE1 driver with numpy standing in for the kernels; X, Y are this rank's rows."""
    K, M = X.shape[1], Y.shape[1]
    S = _allreduce(X.T @ Y)                                  # xty pass + allreduce(K*M)
    W = np.zeros((K, A)); P = np.zeros((K, A)); R = np.zeros((K, A)); Q = np.zeros((M, A)); T = np.zeros((X.shape[0], A))
    for a in range(A):
        if M == 1:
            w = S[:, 0].copy()
        else:
            ev, V = np.linalg.eigh(S.T @ S)
            w = S @ V[:, np.argmax(np.abs(ev))]
        w /= np.sqrt(w @ w)
        r = w - R[:, :a] @ (P[:, :a].T @ w)
        t = X @ r                                            # fused pass on the local rows ...
        gt = _allreduce(np.concatenate([X.T @ t, [t @ t]]))  # ... + allreduce(K+1)
        g, tt = gt[:-1], gt[-1]
        p = g / tt
        q = S.T @ r / tt
        S = S - tt * np.outer(p, q)
        W[:, a], P[:, a], R[:, a], Q[:, a], T[:, a] = w, p, r, q, t
    return dict(W=W, P=P, R=R, Q=Q, T=T)


def _worker(rank, world, port_no, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        seed, N, K, M, A = 31337, 1001, 40, 3, 6
        X, Y = port.synth(seed, N, K, M)
        r0, r1 = sharding.row_shard(N, rank, world)
        st = _sharded_fit(X[r0:r1], Y[r0:r1], A)
        want = port.fit(X, Y, A, 0)
        Bg = port.coefficients(st["R"], st["Q"], A); Bw = port.coefficients(want["R"], want["Q"], A)
        ok_fit = relerr(Bg, Bw) <= RTOL
        sg = np.sign((st["T"] * want["T"][r0:r1]).sum(axis=0))
        ok_T = relerr(st["T"] * sg, want["T"][r0:r1]) <= 1e-8
        # fold sharding: each rank fills its folds' rows of E; the sum over ranks is the full block
        te = np.arange(N, dtype=np.int64).reshape(-1, 1)[:64]
        f0, f1 = sharding.fold_shard(te.shape[0], rank, world)
        full = port.cv_folds(X, Y, A, 0, te)
        mine = np.zeros_like(full)
        mine[:, f0:f1, :] = port.cv_folds(X, Y, A, 0, te[f0:f1])
        ok_cv = np.array_equal(_allreduce(mine), full)
        q.put((rank, bool(ok_fit), bool(ok_T), bool(ok_cv)))
    finally:
        dist.destroy_process_group()


def test_row_and_fold_sharding_world2_gloo():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port_no = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port_no, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted(q.get(timeout=180) for _ in procs)
    [p.join(30) for p in procs]
    assert res == [(0, True, True, True), (1, True, True, True)], res
