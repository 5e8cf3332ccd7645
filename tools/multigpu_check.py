"""Multi-GPU consistency check (run with torchrun, one rank per GPU):

  torchrun --nproc-per-node 2 --master-addr 127.0.0.1 tools/multigpu_check.py

Every rank holds a contiguous shard of the same synthetic data; the results must equal a
single-process run on the whole data: noiseless q-means labels bit-exact and centres to 1e-10,
qPCA singular values to 1e-10, centroid tomography identical on every rank."""
import os
import sys
import warnings

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "sq-learn_b200")):
    sys.path.insert(0, p)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    from oracle import qmeans as oq
    from oracle import qpca as oqp
    from sqlearn.QuantumUtility import Utility as U
    from sqlearn.cluster import qMeans_
    from sqlearn.decomposition import qPCA
    rs = np.random.RandomState(0)
    n, d, k = 40000, 64, 16
    cen = rs.randn(k, d) * 3
    X = cen[rs.randint(0, k, n)] + rs.randn(n, d)
    C0 = X[rs.permutation(n)[:k]].copy()
    lo, hi = n * rank // world, n * (rank + 1) // world
    ok = True
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        U.seed(5)
        est = qMeans_(n_clusters=k, init=C0, n_init=1, max_iter=6, tol=0, delta=0, algorithm='full').fit(X[lo:hi])
        Xc = X - X.mean(axis=0)
        want = oq.kmeans_single_lloyd(Xc, C0 - X.mean(axis=0), False, True, True, True, max_iter=6, tol=0, delta=0,
                                      rng=np.random.default_rng(0))
        ok &= est.labels_ == [int(v) for v in want[0][lo:hi]]
        ok &= np.allclose(est.cluster_centers_, want[2] + X.mean(axis=0), rtol=1e-10, atol=1e-11)
        ok &= abs(est.inertia_ - want[1]) <= 1e-10 * want[1]
        # noisy run: all ranks must hold the same centres (replicated draws agree)
        U.seed(6)
        e2 = qMeans_(n_clusters=k, init=C0, n_init=1, max_iter=3, tol=0, delta=0.5, algorithm='full',
                     intermediate_error=True, true_distance_estimate=False).fit(X[lo:hi])
        c = torch.as_tensor(e2.cluster_centers_).cuda()
        gathered = [torch.empty_like(c) for _ in range(world)]
        dist.all_gather(gathered, c)
        ok &= all(torch.equal(g, gathered[0]) for g in gathered)
        # default seeding (k-means++ on the device, rank 0's centres broadcast): same centres on every rank
        e3 = qMeans_(n_clusters=k, n_init=2, max_iter=4, delta=0.3, algorithm='full', true_distance_estimate=False,
                     random_state=3).fit(X[lo:hi])
        c = torch.as_tensor(e3.cluster_centers_).cuda()
        gathered = [torch.empty_like(c) for _ in range(world)]
        dist.all_gather(gathered, c)
        ok &= all(torch.equal(g, gathered[0]) for g in gathered) and bool(np.isfinite(e3.inertia_))
        pca = qPCA(n_components=8, svd_solver='full').fit(X[lo:hi])
        w = oqp.fit_full(X, 8)
        ok &= np.allclose(pca.singular_values_, w["singular_values_"], rtol=1e-10)
        ok &= np.allclose(pca.components_, w["components_"], atol=1e-8)
        ok &= abs(pca.muA - w["muA"]) <= 1e-10 * w["muA"]
        # one long vector over all ranks (leaf blocks sharded) and a batch of short ones (dealt by vector):
        # bit-identical to the single-GPU routine with the same streams
        for shape, delta, inc in [((1, 300000), 0.05, False), ((2, 70000), 0.3, True), ((37, 500), 0.2, True)]:
            g = torch.Generator(device="cuda"); g.manual_seed(99)
            V = torch.rand(shape, generator=g, device="cuda", dtype=torch.float64) * 2 - 1
            V /= V.norm(dim=1, keepdim=True)
            for norm in ("L2", "inf"):
                U.seed(11)
                got = U.tomography_sharded(V, delta, norm=norm, incremental_measure=inc)
                U.seed(11)
                ref = U._tomography_device(V, delta, None, norm, True, inc, 0)[0]
                same = bool(torch.equal(got, ref))
                if not same:
                    print("rank %d: tomography_sharded differs for %s %s: max |diff| %g" % (
                        rank, shape, norm, float((got - ref).abs().max())), flush=True)
                ok &= same
                lim = delta if norm == "L2" else delta
                e = (got - V).norm(dim=1) if norm == "L2" else (got - V).abs().amax(dim=1)
                ok &= bool((e <= lim).all()) if inc else True
    flag = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("multigpu_check world=%d: %s (q-means labels/centres/inertia, replicated noisy centres, qPCA sigma/V/mu, sharded tomography bit-identical)"
              % (world, "OK" if flag.item() == 1 else "MISMATCH"))
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
