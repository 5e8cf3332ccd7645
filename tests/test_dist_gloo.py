"""N > 1 host logic on CPU: two gloo ranks.  Covers what the multi-GPU path adds on the
host side - the shared Philox state, the one-buffer exchange [k*d sums | k counts | inertia]
and the shard bookkeeping - with the oracle standing in for the kernels."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "sq-learn_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import qmeans as oq
    from sqlearn import _device as dv
    from sqlearn.QuantumUtility import Utility as U

    # 1. every rank ends up with rank 0's seed and stream offset
    U.seed(1000 + rank)
    U._take_streams(17 * (rank + 1))
    U._sync_state()
    state = (U._state["seed"], U._state["offset"])
    assert dv.dist_info() == (rank, world)

    # 2. sharded Lloyd step = single-process Lloyd step
    rs = np.random.RandomState(0)
    n, d, k = 600, 5, 4
    X = rs.randn(n, d) + 3 * rs.randint(0, k, n)[:, None]
    C = X[:k].copy()
    lo, hi = n * rank // world, n * (rank + 1) // world
    Xs = X[lo:hi]
    labels, D, inertia = oq.labels_estimation(Xs, C, 0, True, np.random.default_rng(0))
    red = torch.zeros(k * d + k + 1, dtype=torch.float64)          # same layout as cluster/_engine.py
    sums = red[:k * d].view(k, d)
    counts = red[k * d:k * d + k]
    np.add.at(sums.numpy(), labels, Xs)
    counts += torch.as_tensor(np.bincount(labels, minlength=k), dtype=torch.float64)
    red[-1] = inertia
    dv.all_reduce_sum(red)
    centres = (sums / counts[:, None]).numpy()
    # shard bookkeeping as QMeansEngine does it
    mine = torch.tensor([float(hi - lo)], dtype=torch.float64)
    sizes = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(sizes, mine)
    sizes = [int(s.item()) for s in sizes]
    row0, total = sum(sizes[:rank]), sum(sizes)
    np.save(os.path.join(out_dir, f"r{rank}.npy"),
            np.concatenate([centres.ravel(), counts.numpy(), [float(red[-1]), row0, total, state[0] % 2 ** 31, state[1]]]))
    dist.destroy_process_group()


def test_two_rank_exchange_matches_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    from oracle import qmeans as oq
    rs = np.random.RandomState(0)
    n, d, k = 600, 5, 4
    X = rs.randn(n, d) + 3 * rs.randint(0, k, n)[:, None]
    C = X[:k].copy()
    labels, _, inertia = oq.labels_estimation(X, C, 0, True, np.random.default_rng(0))
    want = oq.centers_update(X, labels, 0, False, True, True, np.random.default_rng(0), n_clusters=k, centers_old=C)
    outs = [np.load(tmp_path / f"r{r}.npy") for r in range(world)]
    for r, o in enumerate(outs):
        np.testing.assert_allclose(o[:k * d].reshape(k, d), want, rtol=1e-12)
        assert np.array_equal(o[k * d:k * d + k], np.bincount(labels, minlength=k))
        assert o[k * d + k] == pytest.approx(inertia, rel=1e-12)
        assert o[k * d + k + 1] == (0 if r == 0 else 300) and o[k * d + k + 2] == 600
    # identical results and identical RNG state on every rank
    assert np.array_equal(outs[0][:k * d + k + 1], outs[1][:k * d + k + 1])
    assert outs[0][-2] == outs[1][-2] == 1000 % 2 ** 31 and outs[0][-1] == outs[1][-1] == 17
