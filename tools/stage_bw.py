"""Bandwidth of the pageable -> device upload path of sqlearn._device.stream_to_device for several host thread
counts (SQ_STAGE_THREADS): usage: python tools/stage_bw.py"""
import os, sys, time, subprocess
if len(sys.argv) > 1:
    import numpy as np, torch
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "sq-learn_b200"))
    from sqlearn import _device as dv
    X = np.random.rand(4_000_000, 128)
    dv.stream_to_device(X[:100000]); torch.cuda.synchronize()
    for rep in range(2):
        t = time.perf_counter(); out = dv.stream_to_device(X); torch.cuda.synchronize(); dt = time.perf_counter() - t
    print("SQ_STAGE_THREADS=%s OMP=%s: %.2f GB/s (%.2f s for %.1f GB)" % (os.environ.get("SQ_STAGE_THREADS"), os.environ.get("OMP_NUM_THREADS"), X.nbytes / 1e9 / dt, dt, X.nbytes / 1e9))
else:
    for thr in ("1", "2", "4", "8", "16"):
        env = dict(os.environ, SQ_STAGE_THREADS=thr, OMP_NUM_THREADS="1")
        subprocess.run([sys.executable, __file__, "x"], env=env)
