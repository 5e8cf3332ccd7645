"""cProfile of qMeans_.fit on rank 0 under torchrun (rows per rank from argv[1])."""
import cProfile, os, pstats, sys, time, warnings
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
import bench
from sqlearn.cluster import qMeans_
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
rows = int(sys.argv[1])
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
X = bench.make_shard(rows * world, rank, world, dev)
Xh = torch.empty(X.shape, dtype=torch.float64, pin_memory=True); Xh.copy_(X)
init = X[:256].cpu().numpy(); del X
init_t = torch.as_tensor(init).cuda(); dist.broadcast(init_t, 0); init = init_t.cpu().numpy()
warnings.simplefilter("ignore")
def fit():
    return qMeans_(n_clusters=256, init=init, n_init=1, max_iter=10, tol=0, delta=0.5, algorithm='full',
                   intermediate_error=True, true_distance_estimate=False, copy_x=False).fit(Xh.numpy())
fit(); dist.barrier(); torch.cuda.synchronize()
pr = cProfile.Profile()
t = time.perf_counter(); pr.enable(); fit(); torch.cuda.synchronize(); pr.disable(); dt = time.perf_counter() - t
if rank == 0:
    print("fit %.1f ms at world %d, %d rows per rank" % (dt * 1e3, world, rows))
    pstats.Stats(pr).sort_stats("tottime").print_stats(18)
dist.destroy_process_group()
