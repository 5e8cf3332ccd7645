"""Quantum routines of sq-learn (tomography, amplitude / phase estimation, ...)
on a B200.

Host-side mirror of the reference module ``sklearn/QuantumUtility/Utility.py``:
same function names, argument meaning and error behaviour, so code written
against ``sklearn.QuantumUtility`` runs unchanged against
``sqlearn.QuantumUtility``.  The sampling itself happens in hand-written CUDA
kernels reached through the C ABI (``include/sqb200.h``); this file only
validates arguments, computes the integer schedules exactly as the reference
does, and moves arrays.  There is no CPU fallback.

Randomness: the reference draws from OS entropy / Python's global ``random``
(Utility.py:53, :508, :653), so its results are not reproducible; here every
call consumes a fresh range of counter-based Philox streams from a module
state that :func:`seed` resets.  Noisy results therefore match the reference
in distribution, never bit for bit.
"""
import math
import os
import warnings
from collections import Counter

import ctypes

import numpy as np
import torch

from .. import _cabi
from .. import _device as dv

warnings.simplefilter('always', UserWarning)  # as the reference does (Utility.py:18)

# --------------------------------------------------------------------------
# random state
# --------------------------------------------------------------------------
_state = {"seed": int.from_bytes(os.urandom(8), "little"), "offset": 0}


def seed(s):
    """Make every later draw of this module reproducible."""
    _state["seed"] = int(s) & 0xFFFFFFFFFFFFFFFF
    _state["offset"] = 0


def _take_streams(n):
    """Reserve ``n`` Philox streams; returns (seed, first offset)."""
    off = _state["offset"]
    _state["offset"] = off + int(n)
    return _state["seed"], off


# --------------------------------------------------------------------------
# small classical helpers kept for API compatibility
# --------------------------------------------------------------------------
class QuantumState(object):
    """A register whose outcome i appears with probability amplitude_i^2
    (reference: Utility.py:25-58).  ``measure`` draws on the GPU."""

    def __init__(self, registers, amplitudes):
        self.registers = registers
        amps = np.asarray(amplitudes, dtype=np.float64)
        self.norm_factor = math.sqrt(float(np.sum(amps * amps)))
        self.amplitudes = list(amps / self.norm_factor)
        self.probabilities = list((amps / self.norm_factor) ** 2)
        assert len(self.registers) == len(self.amplitudes)
        assert abs(sum(self.probabilities) - 1) < 0.0000000001

    def measure(self, n_times=1):
        counts = multinomial_counts(np.asarray(self.probabilities), int(n_times))
        out = np.repeat(np.asarray(self.registers), counts.astype(np.int64))
        # the reference returns the samples in draw order; a multiset in random order is the same law
        perm = np.random.default_rng(_take_streams(1)[1] ^ _state["seed"]).permutation(len(out))
        return out[perm]

    def get_state(self):
        return {self.registers[i]: self.probabilities[i] for i in range(len(self.registers))}


def estimate_wald(measurements):
    """Relative frequencies of the outcomes (Utility.py:61-64)."""
    counter = Counter(measurements)
    return {x: counter[x] / len(measurements) for x in counter}


def multinomial_counts(p, N):
    """Counts of N draws from the distribution p, drawn on the GPU by the
    binomial splitting the tomography kernel uses (one level at a time)."""
    p = np.asarray(p, dtype=np.float64)
    d = len(p)
    L = 1
    while L < d:
        L *= 2
    mass = np.zeros(2 * L)
    mass[L:L + d] = p
    for i in range(L - 1, 0, -1):
        mass[i] = mass[2 * i] + mass[2 * i + 1]
    cnt = np.zeros(2 * L)
    cnt[1] = N
    first = 1
    while first < L:
        idx = np.arange(first, 2 * first)
        left = _binomial(cnt[idx], mass[2 * idx], mass[2 * idx + 1])
        cnt[2 * idx] = left
        cnt[2 * idx + 1] = cnt[idx] - left
        first *= 2
    return cnt[L:L + d]


def _binomial(n, pl, pr):
    n = dv.to_device(np.asarray(n, dtype=np.float64))
    pl = dv.to_device(np.asarray(pl, dtype=np.float64))
    pr = dv.to_device(np.asarray(pr, dtype=np.float64))
    out = torch.empty_like(n)
    s, off = _take_streams(n.numel())
    _cabi.check(_cabi.lib.sq_binomial_sample(dv.ptr(n), dv.ptr(pl), dv.ptr(pr), n.numel(), s, off, dv.ptr(out),
                                             dv.stream_ptr()), "sq_binomial_sample")
    return out.cpu().numpy()


def coupon_collect(quantum_state):
    """Number of measurements until every outcome was seen (Utility.py:75-85)."""
    counter = 0
    collection = {value: 0 for value in quantum_state.get_state().keys()}
    while sum(collection.values()) != len(collection):
        value = quantum_state.measure()[0]
        if not collection[value]:
            collection[value] = 1
        counter += 1
    return counter


def create_rand_vec(n_vec, len_vec, scale=None, type='uniform'):
    """Random unit vectors (Utility.py:183-193); host NumPy, as in the reference."""
    v = []
    for _ in range(n_vec):
        if type == 'uniform':
            vv = np.random.uniform(-1, 1, (len_vec))
        elif type == 'exp':
            vv = np.random.exponential(scale=scale, size=len_vec)
        vv = vv / np.linalg.norm(vv, ord=2)
        v.append(vv)
    return v


def check_measure(arr, faster_measure_increment):
    """Make the attempt schedule strictly increasing (Utility.py:414-422)."""
    incr = 5 + faster_measure_increment
    for i in range(len(arr) - 1):
        if arr[i + 1] == arr[i]:
            arr[i + 1] += incr
        if arr[i + 1] <= arr[i]:
            arr[i + 1] = arr[i] + incr
    return arr


def check_division(v, n_jobs):
    """Split v work items over n_jobs workers (Utility.py:425-432)."""
    a = float(v) / n_jobs
    frac = a - int(a)
    remaining = int(frac * n_jobs)
    out = [int(a) for _ in range(n_jobs)]
    for i in range(remaining):
        out[i] += 1
    return out


def amplitude_est_dist(w0, w1):
    """Circular distance between two phases in units of a turn (Utility.py:435-439)."""
    c = -np.ceil(w1 - w0)
    f = -np.floor(w1 - w0)
    return min(np.abs(c + w1 - w0), np.abs(f + w1 - w0))


# --------------------------------------------------------------------------
# truncated-normal error injection (K8)
# --------------------------------------------------------------------------
def _truncnorm_add(x, halfwidth):
    """x + TruncNormal(0, 1; [-halfwidth, halfwidth]) element-wise on the GPU."""
    is_t = isinstance(x, torch.Tensor)
    xd = dv.to_device(x) if not is_t else x.contiguous()
    out = torch.empty_like(xd)
    s, off = _take_streams(xd.numel())
    _cabi.check(_cabi.lib.sq_truncnorm_add(dv.ptr(xd), xd.numel(), float(halfwidth), s, off, dv.ptr(out),
                                           dv.stream_ptr()), "sq_truncnorm_add")
    return out if is_t else out.cpu().numpy()


def introduce_error(value, epsilon):
    """value + truncnorm(-epsilon, epsilon) (Utility.py:68-69); returns shape (1,)."""
    return _truncnorm_add(np.asarray([value], dtype=np.float64), epsilon)


def introduce_error_array(array, norm_error):
    """array + truncnorm(+-norm_error/sqrt(size)) per entry (Utility.py:71-73)."""
    size = array.shape[0]
    return _truncnorm_add(array, norm_error / np.sqrt(size))


def make_gaussian_est(vec, noise):
    """Gaussian-noise approximation of tomography (Utility.py:88-104)."""
    noise_per_component = noise / np.sqrt(len(vec))
    if noise_per_component != 0:
        new_vec = _truncnorm_add(vec, noise_per_component)
    else:
        # the reference leaves new_vec unbound here (UnboundLocalError)
        raise UnboundLocalError("local variable 'new_vec' referenced before assignment")
    return new_vec


# --------------------------------------------------------------------------
# tomography (K1)
# --------------------------------------------------------------------------
def tomography_num_measurements(d, delta, norm='L2'):
    """N of real_tomography (Utility.py:307-311)."""
    if norm == 'L2':
        return int((36 * d * np.log(d)) / (delta ** 2))
    elif norm == 'inf':
        return int((36 * np.log(d)) / (delta ** 2))
    raise ValueError("norm must be 'L2' or 'inf'")


def tomography_schedule(N, faster_measure_increment=0):
    """Attempt sizes of the incremental mode (Utility.py:317-319)."""
    idx = np.geomspace(1, N, num=100, dtype=np.int64)
    return check_measure(idx, faster_measure_increment=faster_measure_increment)


_spec_states = {}


def _spec_state(B):
    """Search state of sq_tomography_batched_spec_fused for batches of B vectors on the current device and
    stream: prepared once, left prepared by every call."""
    key = (torch.cuda.current_device(), torch.cuda.current_stream().cuda_stream, int(B))
    st = _spec_states.get(key)
    if st is None:
        st = dv.empty((2 * int(B),), dtype=torch.int32)
        _cabi.check(_cabi.lib.sq_tomography_spec_state_init(dv.ptr(st), int(B), dv.stream_ptr()),
                    "sq_tomography_spec_state_init")
        if len(_spec_states) > 64:
            _spec_states.clear()
        _spec_states[key] = st
    return st


def _tomography_device(V, delta, N, norm, stop, incremental, fmi, trace=False, vec_offset=None, seed_=None):
    """Run K1 on a device tensor V [B, d]; returns (est, n_used, err, trace, attempts)."""
    B, d = V.shape
    if N is None:
        N = tomography_num_measurements(d, delta, norm)
    if incremental:
        sched = tomography_schedule(int(N), fmi).astype(np.float64)
        if not stop:
            # every attempt is a fresh draw and only the last one is returned unless a trace is wanted
            if not trace:
                sched = sched[-1:]
    else:
        sched = np.asarray([float(int(N))])
    sched = np.ascontiguousarray(sched, dtype=np.float64)
    est = torch.empty_like(V)
    n_used = dv.empty((B,))
    err = dv.empty((B,))
    if B == 0:
        return est, n_used, err, None, None, sched
    tr = dv.empty((B, len(sched), d)) if trace else None
    att = dv.empty((B,), dtype=torch.int32) if trace else None
    if seed_ is None:
        seed_, vec_offset = _take_streams(B)
    if d > _cabi.lib.sq_tomography_max_fused_dim():
        if trace:
            raise NotImplementedError("the per-attempt trace is kept only for d <= %d"
                                      % _cabi.lib.sq_tomography_max_fused_dim())
        # long vectors: grid-wide path with caller workspace; at most 65535 vectors per call
        for b0 in range(0, B, 65535):
            b1 = min(B, b0 + 65535)
            nb = _cabi.lib.sq_tomography_workspace_bytes(b1 - b0, d)
            ws = dv.workspace(nb)
            _cabi.check(_cabi.lib.sq_tomography_batched_ws(
                dv.ptr(V[b0:b1]), b1 - b0, d, float(delta), 1 if norm == 'inf' else 0,
                1 if (stop and incremental) else 0, sched.ctypes.data, len(sched), seed_, vec_offset + b0,
                dv.ptr(est[b0:b1]), dv.ptr(n_used[b0:b1]), dv.ptr(err[b0:b1]), dv.ptr(ws), nb, dv.stream_ptr()),
                "sq_tomography_batched_ws")
        return est, n_used, err, tr, att, sched
    if stop and incremental and len(sched) > 1 and not trace and B <= 2048:
        # few vectors: the latency of the attempt-by-attempt search dominates; 8 attempt lanes per vector search
        # the schedule together and the lane that finishes last replays the first successful attempt (one launch)
        lanes = 8        # (any number of lanes gives the same bits; 8 measured fastest from 16 to 256 vectors)
        _cabi.check(_cabi.lib.sq_tomography_batched_spec_fused(
            dv.ptr(V), B, d, float(delta), 1 if norm == 'inf' else 0, sched.ctypes.data, len(sched), lanes, seed_,
            vec_offset, dv.ptr(est), dv.ptr(n_used), dv.ptr(err), None, dv.ptr(_spec_state(B)), dv.stream_ptr()),
            "sq_tomography_batched_spec_fused")
        return est, n_used, err, tr, att, sched
    if not trace:
        # a single attempt of many vectors runs in the group kernel, which keeps its trees in a workspace
        # (0 bytes asked for when the fused kernel will be used)
        nb = _cabi.lib.sq_tomography_workspace_bytes(B, d)
        ws = dv.workspace(nb) if nb else None
        _cabi.check(_cabi.lib.sq_tomography_batched_ws(
            dv.ptr(V), B, d, float(delta), 1 if norm == 'inf' else 0, 1 if (stop and incremental) else 0,
            sched.ctypes.data, len(sched), seed_, vec_offset, dv.ptr(est), dv.ptr(n_used), dv.ptr(err),
            dv.ptr(ws), nb, dv.stream_ptr()), "sq_tomography_batched_ws")
        return est, n_used, err, tr, att, sched
    _cabi.check(_cabi.lib.sq_tomography_batched(
        dv.ptr(V), B, d, float(delta), 1 if norm == 'inf' else 0, 1 if (stop and incremental) else 0,
        sched.ctypes.data, len(sched), seed_, vec_offset, dv.ptr(est), dv.ptr(n_used), dv.ptr(err),
        dv.ptr(tr), dv.ptr(att), dv.stream_ptr()), "sq_tomography_batched")
    return est, n_used, err, tr, att, sched


def tomography_sharded(A, delta, N=None, norm='L2', stop_when_reached_accuracy=True, incremental_measure=True,
                       faster_measure_increment=0):
    """``tomography(A, delta)`` computed by all the ranks of the process group together (collective: every rank
    passes the same ``A``, [d] or [B, d], and gets the same estimate back as a device tensor).

    Short vectors (d <= sq_tomography_max_fused_dim()) are dealt to the ranks by vector.  A long vector is cut
    into blocks of leaves and rank r splits the measurement counts of its share of the blocks (reference: the
    loop of real_tomography, Utility.py:259-402); per attempt the ranks exchange two small sums
    (block masses, block error partials).  Draws are keyed by (vector, block, node), so the estimate is
    bit-identical to the single-GPU one for any number of ranks."""
    import torch.distributed as dist
    if norm not in ('L2', 'inf'):
        raise ValueError("norm must be 'L2' or 'inf'")
    V = A if isinstance(A, torch.Tensor) else dv.to_device(np.asarray(A, dtype=np.float64))
    one = V.dim() == 1
    V = V.reshape(1, -1) if one else V
    V = V.contiguous()
    B, d = V.shape
    rank, world = dv.dist_info()
    _sync_state()
    stop = bool(stop_when_reached_accuracy)
    if world == 1 or B == 0:
        est = _tomography_device(V, delta, N, norm, stop, incremental_measure, faster_measure_increment)[0]
        return est[0] if one else est
    seed_, off = _take_streams(B)
    est = torch.zeros_like(V)
    if d <= _cabi.lib.sq_tomography_max_fused_dim():
        lo, hi = (B * rank) // world, (B * (rank + 1)) // world
        if hi > lo:
            est[lo:hi] = _tomography_device(V[lo:hi], delta, N, norm, stop, incremental_measure,
                                            faster_measure_increment, vec_offset=off + lo, seed_=seed_)[0]
        dist.all_reduce(est)                       # every row is written by one rank, zero elsewhere
        return est[0] if one else est
    if B > 65535:
        raise ValueError("tomography_sharded: at most 65535 long vectors per call")
    if N is None:
        N = tomography_num_measurements(d, delta, norm)
    sched = tomography_schedule(int(N), faster_measure_increment).astype(np.float64) if incremental_measure \
        else np.asarray([float(int(N))])
    if incremental_measure and not stop:
        sched = sched[-1:]
    lay = (ctypes.c_longlong * 8)()
    _cabi.check(_cabi.lib.sq_tomography_workspace_layout(B, d, lay), "sq_tomography_workspace_layout")
    nblk, tb, nb = int(lay[0]), int(lay[1]), int(lay[7])
    ws = dv.workspace(nb)

    def view(i, dtype=torch.float64, cols=nblk):
        o = int(lay[2 + i])
        return ws[o:o + B * cols * (8 if dtype == torch.float64 else 4)].view(dtype).view(B, cols)
    bm = ws[int(lay[2]):int(lay[4])].view(torch.float64)      # bm_a | bm_b, adjacent in the layout
    bm_a, bm_b, errp = view(0), view(1), view(3)
    done = ws[int(lay[6]):int(lay[6]) + 4 * B].view(torch.int32)
    blo, bhi = (nblk * rank) // world, (nblk * (rank + 1)) // world
    n_used, err = dv.empty((B,)), dv.empty((B,))

    def phase(ph, at, n_meas):
        _cabi.check(_cabi.lib.sq_tomography_sharded_phase(
            dv.ptr(V), B, d, float(delta), 1 if norm == 'inf' else 0, 1 if (stop and incremental_measure) else 0,
            float(n_meas), at, seed_, off, dv.ptr(est), dv.ptr(n_used), dv.ptr(err), dv.ptr(ws), nb, blo, bhi, ph,
            dv.stream_ptr()), "sq_tomography_sharded_phase")
    phase(0, 0, 1.0)
    for at, n_meas in enumerate(sched):
        phase(1, at, n_meas)
        bm_a[:, :blo] = 0; bm_a[:, bhi:] = 0; bm_b[:, :blo] = 0; bm_b[:, bhi:] = 0
        dist.all_reduce(bm)
        phase(2, at, n_meas)
        errp[:, :blo] = 0; errp[:, bhi:] = 0
        dist.all_reduce(errp, op=dist.ReduceOp.MAX if norm == 'inf' else dist.ReduceOp.SUM)
        phase(3, at, n_meas)
        if stop and incremental_measure and at + 1 < len(sched) and bool(done.all()):
            break
    est[:, :blo * tb] = 0
    est[:, bhi * tb:] = 0
    dist.all_reduce(est)
    return est[0] if one else est


def real_tomography(V, N=None, delta=None, stop_when_reached_accuracy=True, norm='L2', incremental_measure=True,
                    faster_measure_increment=0):
    """Vector-state tomography of one vector (reference: Utility.py:259-402).

    Returns ``{N_measure: estimate}`` with one entry per attempt made, like
    the reference; the last entry is the estimate ``tomography`` returns.
    """
    if norm not in ('L2', 'inf'):
        raise ValueError("norm must be 'L2' or 'inf'")
    Vd = dv.to_device(np.asarray(V, dtype=np.float64).reshape(1, -1))
    if Vd.shape[1] > _cabi.lib.sq_tomography_max_fused_dim():
        # long vectors: only the returned attempt is kept (its estimate alone is d doubles per attempt)
        est, n_used, err, _, _, _ = _tomography_device(
            Vd, delta, N, norm, stop_when_reached_accuracy, incremental_measure, faster_measure_increment)
        return {int(n_used.item()): list(est[0].cpu().numpy())}
    est, n_used, err, tr, att, sched = _tomography_device(
        Vd, delta, N, norm, stop_when_reached_accuracy, incremental_measure, faster_measure_increment, trace=True)
    n_att = int(att.item())
    tr = tr[0, :n_att].cpu().numpy()
    return {int(sched[i]): list(tr[i]) for i in range(n_att)}


def tomography(A, noise, true_tomography=True, stop_when_reached_accuracy=True, N=None, norm='L2',
               incremental_measure=True, faster_measure_increment=0):
    """Tomography of a vector or of every row of a matrix (reference: Utility.py:107-180).

    ``A`` may be a NumPy array (a NumPy array is returned) or a CUDA tensor
    (a CUDA tensor is returned; used by the estimators to stay on the device).
    """
    assert noise >= 0
    if noise == 0:
        return A
    is_t = isinstance(A, torch.Tensor)
    shape = tuple(A.shape)
    if true_tomography == False:  # noqa: E712  (the reference's comparison)
        if len(shape) == 2:
            flat = A.reshape(shape[0] * shape[1])
            return make_gaussian_est(flat, noise).reshape(shape[0], shape[1])
        return make_gaussian_est(A, noise)
    if norm not in ('L2', 'inf'):
        raise ValueError("norm must be 'L2' or 'inf'")
    Ad = A if is_t else dv.to_device(np.asarray(A, dtype=np.float64))
    Ad = Ad.reshape(-1, shape[-1]).contiguous() if len(shape) == 2 else Ad.reshape(1, -1).contiguous()
    est = _tomography_device(Ad, noise, N, norm, stop_when_reached_accuracy, incremental_measure,
                             faster_measure_increment)[0]
    est = est.reshape(shape)
    return est if is_t else est.cpu().numpy()


def L2_tomogrphy_fakeSign(V, N=None, delta=None):
    """First half of the tomography with the true signs copied (Utility.py:234-256)."""
    V = np.asarray(V, dtype=np.float64)
    d = len(V)
    if N is None:
        N = (36 * d * np.log(d)) / (delta ** 2)
    p = V ** 2 / np.sum(V ** 2)
    counts = multinomial_counts(p, int(N))
    est = np.sqrt(counts / int(N))
    return list(np.where(V < 0, -est, est))


# --------------------------------------------------------------------------
# mu-norm search (K6)
# --------------------------------------------------------------------------
def _mu_grid_device(X, ps):
    """sqrt(s1*s2) for every p of the grid on a device matrix X (rows = this rank's shard)."""
    n, d = X.shape
    ps = [float(p) for p in ps]
    er = np.asarray([2 * p for p in ps], dtype=np.float64)
    ec = np.asarray([2 * (1 - p) for p in ps], dtype=np.float64)
    vals = np.zeros(len(ps))
    frob2 = 0.0
    for c0 in range(0, len(ps), 16):
        c1 = min(c0 + 16, len(ps))
        npk = c1 - c0
        rowmax = dv.empty((npk,))
        colsum = dv.empty((npk, d))
        fr = dv.empty((1,))
        nbytes = _cabi.lib.sq_mu_norms_workspace_bytes(n, d, npk)
        ws = dv.workspace(nbytes)
        e_r = np.ascontiguousarray(er[c0:c1])
        e_c = np.ascontiguousarray(ec[c0:c1])
        _cabi.check(_cabi.lib.sq_mu_norms(dv.ptr(X), n, d, e_r.ctypes.data, e_c.ctypes.data, npk, dv.ptr(rowmax),
                                          dv.ptr(colsum), dv.ptr(fr), dv.ptr(ws), nbytes, dv.stream_ptr()),
                    "sq_mu_norms")
        rank, world = dv.dist_info()
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(rowmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(colsum, op=dist.ReduceOp.SUM)
            dist.all_reduce(fr, op=dist.ReduceOp.SUM)
        s1 = rowmax.cpu().numpy()
        s2 = colsum.max(dim=1).values.cpu().numpy()
        vals[c0:c1] = np.sqrt(s1 * s2)
        frob2 = float(fr.item())
    return vals, math.sqrt(frob2)


def linear_search(matrix, start=0.0, end=1.0, step=0.05):
    """Best p on the grid and its mu value (Utility.py:215-219)."""
    domain = [i for i in np.arange(start, end, step)] + [end]
    X = matrix if isinstance(matrix, torch.Tensor) else dv.to_device(np.asarray(matrix, dtype=np.float64))
    values, _ = _mu_grid_device(X, domain)
    values = list(values)
    best_p = domain[values.index(min(values))]
    return best_p, min(values)


def best_mu(matrix, start=0.0, end=1.0, step=0.05):
    """min(mu_p over the p grid, Frobenius norm) with its name (Utility.py:222-231)."""
    domain = [i for i in np.arange(start, end, step)] + [end]
    X = matrix if isinstance(matrix, torch.Tensor) else dv.to_device(np.asarray(matrix, dtype=np.float64))
    values, frob = _mu_grid_device(X, domain)
    values = list(values)
    val = min(values)
    p = domain[values.index(val)]
    val_list = [val, frob]
    index = val_list.index(min(val_list))
    if index == 0:
        best_norm = f"p={p}"
    else:
        best_norm = "Frobenius"
    return best_norm, val_list[index]


# --------------------------------------------------------------------------
# amplitude / phase estimation (K7)
# --------------------------------------------------------------------------
def _fejer_draw(M, x, brute=False):
    """One outcome j per (M, x) pair; arrays in, float64 array out."""
    M = np.ascontiguousarray(np.broadcast_to(np.asarray(M, dtype=np.float64), np.shape(x)).ravel())
    x = np.ascontiguousarray(np.asarray(x, dtype=np.float64).ravel())
    Md, xd = dv.to_device(M), dv.to_device(x)
    out = torch.empty_like(xd)
    s, off = _take_streams(xd.numel())
    fn = _cabi.lib.sq_fejer_sample_bruteforce if brute else _cabi.lib.sq_fejer_sample
    _cabi.check(fn(dv.ptr(Md), dv.ptr(xd), xd.numel(), s, off, dv.ptr(out), dv.stream_ptr()), "sq_fejer_sample")
    return out.cpu().numpy()


def median_evaluation_Q(gamma):
    """Number of repetitions of median_evaluation (Utility.py:564-568)."""
    z = np.log(1 / gamma) / (2 * (8 / np.pi ** 2 - 0.5) ** 2)
    Q = np.ceil(z)
    if Q % 2 == 0:
        Q += 1
    return int(Q)


def median_evaluation(func, gamma=0.1, Q=None, *args, **kwargs):
    """Median of Q runs of ``func`` (Utility.py:534-572)."""
    if Q == None:  # noqa: E711
        Q = median_evaluation_Q(gamma)
    estimates = [func(*args, **kwargs) for _ in range(int(Q))]
    return np.median(estimates)


def amplitude_estimation_M(epsilon):
    """Grid size of amplitude estimation (Utility.py:484)."""
    return math.ceil((np.pi / (2 * epsilon)) * (1 + np.sqrt(1 + 4 * epsilon)))


def amplitude_estimation(a, epsilon=0.01, gamma=None, M=None, nqubit=False, plot_distribution=False):
    """Amplitude estimation of a in [0,1] (reference: Utility.py:442-531)."""
    if plot_distribution:
        warnings.warn("plot_distribution is not supported in this build (no matplotlib); ignored")
    if M == None:  # noqa: E711
        M = amplitude_estimation_M(epsilon)
        n_qubits = np.ceil(np.log2(M))
    else:
        n_qubits = np.ceil(np.log2(M))
        warnings.warn(
            "Attention! The value of M that will be considered is the one you passed. Epsilon in this case is "
            "useless")
    theta_a = math.asin(np.sqrt(a))
    x = M * (theta_a / np.pi)
    if gamma:
        Q = median_evaluation_Q(gamma)
        j = _fejer_draw(float(M), np.full(Q, x))
        return float(np.median(np.sin(np.pi * j / M) ** 2))
    j = float(_fejer_draw(float(M), np.asarray([x]))[0])
    theta_tilde = np.pi * j / M
    if nqubit:
        return theta_tilde, n_qubits, M
    return np.sin(theta_tilde) ** 2


def wrapper_phase_est_arguments(argument, type='sv'):
    """Map a singular value / distance to a phase (Utility.py:575-581)."""
    if type == 'sv':
        return 2 * math.acos(argument)
    if type == 'distance':
        return math.asin(np.sqrt(argument))


def unwrap_phase_est_arguments(argument, eps, type='sv'):
    """Inverse of wrapper_phase_est_arguments (Utility.py:584-588)."""
    if type == 'sv':
        return math.cos(argument * (eps + np.pi) / 2)
    if type == 'distance':
        return math.sin(argument * np.pi) ** 2


def phase_estimation_m(epsilon, gamma):
    """Number of qubits of phase estimation (Utility.py:635)."""
    return int(np.ceil(np.log2(1 / epsilon)) + np.ceil(np.log2(2 + 1 / (2 * gamma))))


def _phase_estimation_batch(omegas, m):
    """k~/M for an array of phases at M = 2^m (no warnings; used by the estimators)."""
    omegas = np.asarray(omegas, dtype=np.float64)
    M = float(2 ** m)
    out = np.empty_like(omegas)
    one = (omegas == 1) | np.isclose(omegas, 1)
    out[one] = (M - 1) / M
    rest = ~one
    if rest.any():
        k = _fejer_draw(M, omegas[rest] * M)
        out[rest] = k / M
    return out


def phase_estimation(omega, m=None, epsilon=None, gamma=0.1, plot_distribution=False, nqubit=False):
    """Phase estimation of omega in [0,1) (reference: Utility.py:591-694)."""
    assert m != None or epsilon != None, \
        "Attention! You need to specify the number of qubits m or the precision epsilon."  # noqa: E711
    if m != None and nqubit == True:  # noqa: E711,E712
        warnings.warn("Attention! You are specifying that you want to return also the number of qubits"
                      " used, but you are already specifying it with the m parameter.")
    if epsilon != None:  # noqa: E711
        warnings.warn('Attention! The m value is computed using the epsilon parameter passed.')
        m = phase_estimation_m(epsilon, gamma)
    if plot_distribution:
        warnings.warn("plot_distribution is not supported in this build (no matplotlib); ignored")
    M = 2 ** m
    if omega == 1 or np.isclose(omega, 1):
        return (M - 1) / M
    omega_tilde = float(_phase_estimation_batch(np.asarray([omega]), m)[0])
    if nqubit:
        return omega_tilde, omega_tilde * M, m, M
    return omega_tilde


def _consistent_bucket(pe_estimate, epsilon, shift, delta_prime):
    """Midpoint of the grid cell holding pe_estimate (Utility.py:782-790), without
    materialising the np.arange grid: cell edges are start + i*epsilon."""
    start = -1 - shift * delta_prime
    stop = 1 + epsilon - shift * delta_prime
    n_cells = int(np.ceil((stop - start) / epsilon))  # length of np.arange(start, stop, epsilon)
    last = n_cells  # index of the appended edge `stop`
    step = (start + epsilon) - start  # np.arange fills start + i*(second - first)

    def edge(i):
        return stop if i >= last else start + i * step

    i = int(np.floor((pe_estimate - start) / epsilon)) + 1  # bisect_right candidate
    i = max(0, min(i, last + 1))
    while i <= last and edge(i) <= pe_estimate:
        i += 1
    while i > 0 and edge(i - 1) > pe_estimate:
        i -= 1
    if i > last:          # pe_estimate beyond the grid: the reference raises IndexError
        raise IndexError("index %d is out of bounds for axis 0 with size %d" % (i, last + 1))
    estimate = float(np.mean((edge(i - 1), edge(i))))
    return 0 if estimate < 0 else estimate


def _consistent_phase_estimation_batch(omegas, epsilon, gamma):
    """Vectorised consistent phase estimation for the estimators (no warnings)."""
    n = phase_estimation_m(epsilon, gamma)
    C = gamma / n
    delta_prime = (epsilon * C) / 2
    L = np.floor(2 / C)
    shift = int(L / 2) + 1
    m = phase_estimation_m(delta_prime, gamma)
    pe = _phase_estimation_batch(omegas, m)
    return np.asarray([_consistent_bucket(v, epsilon, shift, delta_prime) for v in pe])


def consistent_phase_estimation(omega, epsilon, gamma, n=None, shift=None):
    """Consistent phase estimation (reference: Utility.py:740-792)."""
    if n == None:  # noqa: E711
        n = phase_estimation_m(epsilon, gamma)
    C = gamma / n
    delta_prime = (epsilon * C) / 2
    L = np.floor(2 / C)
    if shift == None:  # noqa: E711
        shift = int(L / 2) + 1
    pe_estimate = phase_estimation(omega=omega, epsilon=delta_prime, gamma=gamma)
    return _consistent_bucket(pe_estimate, epsilon, shift, delta_prime)


def ipe(x, y, epsilon, Q=1, gamma=0.1):
    """Robust inner product estimation (reference: Utility.py:697-737)."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    S = np.linalg.norm(x) ** 2 + np.linalg.norm(y) ** 2
    inner = np.inner(x, y)
    a = (S - 2 * inner) / (2 * S)
    epsilon_a = epsilon * max(1, np.abs(inner)) / S
    if math.isclose(a, 0.0, abs_tol=1e-15):
        a = 0
    a_tilde = amplitude_estimation(a=a, gamma=gamma, epsilon=epsilon_a)
    return S * (1 - 2 * a_tilde) / 2


def _sync_state():
    """Give every rank of the process group rank 0's seed and stream offset, so
    that replicated draws (centroid tomography, singular-value estimation) agree."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return
    dev = dv.require_cuda() if dist.get_backend() == "nccl" else torch.device("cpu")
    s = _state["seed"]
    t = torch.tensor([s & 0xFFFFFFFF, s >> 32, _state["offset"] & 0xFFFFFFFF, _state["offset"] >> 32],
                     dtype=torch.int64, device=dev)
    dist.broadcast(t, src=0)
    v = [int(x) for x in t.cpu().tolist()]
    _state["seed"] = v[0] | (v[1] << 32)
    _state["offset"] = v[2] | (v[3] << 32)
