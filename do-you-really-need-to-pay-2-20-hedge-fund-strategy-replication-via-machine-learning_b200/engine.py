"""Host side of the batched replication path: prepares inputs the way the reference does
(autoencoder_v4.ipynb:112, Autoencoder_encapsulate.py:62-67), moves them to the GPU and drives
the three CUDA kernels through the C ABI (include/hfrep.h).  PyTorch is used for device
memory and streams only.

A *task* is one independent autoencoder fit plus its evaluation (Mode R):
task = (latent_dim L, initial kernels W0[, per-epoch row orders]).
"""
from __future__ import annotations

import ctypes as C
import dataclasses

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, ptr, stream_ptr


# ------------------------------------------------------------------------------------ host prep
def minmax_scale(x: np.ndarray) -> np.ndarray:
    """MinMaxScaler().fit_transform in float64 (Autoencoder_encapsulate.py:62-65): scale = 1/(max-min)
    (1 where the range is zero), out = x*scale + (0 - min*scale)."""
    x = np.asarray(x, dtype=np.float64)
    lo = x.min(axis=0)
    rng = x.max(axis=0) - lo
    rng[rng == 0.] = 1.
    scale = 1. / rng
    return x * scale + (0. - lo * scale)


def split_rows(n: int, test_size: float = .5):
    """train_test_split(shuffle=False, test_size) row counts (autoencoder_v4.ipynb:112): n_test = ceil(test_size*n)."""
    n_test = int(np.ceil(test_size * n))
    return n - n_test, n_test


def glorot_limit(fan_in: int, fan_out: int) -> float:
    return float(np.sqrt(6. / (fan_in + fan_out)))


def initial_weights(seed: int, E: int, L: int):
    """GlorotUniform kernels (Dense default, Autoencoder_encapsulate.py:24,28) from numpy's Philox
    generator: encoder stream keyed 2*seed, decoder stream 2*seed+1, U(-l, l) drawn in float64 and
    rounded to float32.  (Keras draws from TensorFlow's RNG, which is not reproducible here.)"""
    out = []
    for key, (fi, fo) in ((2 * seed, (E, L)), (2 * seed + 1, (L, E))):
        rng = np.random.Generator(np.random.Philox(key=int(key)))
        lim = glorot_limit(fi, fo)
        out.append(rng.uniform(-lim, lim, (fi, fo)).astype(np.float32))
    return out[0], out[1]


def device_permutations(stream_seed: int, epochs: int, n_fit: int) -> np.ndarray:
    """Host replica of the kernel's device-side epoch shuffles (hfrep.h, task_seed): stable argsort of murmur3-finalised
    keys of (seed, epoch, row).  For inspection and tests; the kernel never needs the table."""
    M = np.uint64(0xFFFFFFFF)
    i = np.arange(1, n_fit + 1, dtype=np.uint64)
    out = np.empty((epochs, n_fit), np.int32)
    for e in range(epochs):
        h = (np.uint64(stream_seed) * np.uint64(0x9E3779B1) & M) ^ (np.uint64(e + 1) * np.uint64(0x85EBCA77) & M) ^ (i * np.uint64(0xC2B2AE3D) & M)
        h ^= h >> np.uint64(16); h = h * np.uint64(0x85EBCA6B) & M
        h ^= h >> np.uint64(13); h = h * np.uint64(0xC2B2AE35) & M
        h ^= h >> np.uint64(16)
        out[e] = np.argsort(h, kind="stable")
    return out


def epoch_permutations(seed: int, epochs: int, n_fit: int) -> np.ndarray:
    """Per-epoch row orders standing in for Keras' shuffle=True (int32 [epochs][n_fit])."""
    rng = np.random.Generator(np.random.Philox(key=int(seed)))
    return np.stack([rng.permutation(n_fit) for _ in range(epochs)]).astype(np.int32)


def pack_weights(pairs, E: int, w_stride: int) -> np.ndarray:
    """[(We (E,L), Wd (L,E)), ...] -> f32 [n][w_stride] in the C ABI's packing."""
    out = np.zeros((len(pairs), w_stride), np.float32)
    for i, (We, Wd) in enumerate(pairs):
        L = We.shape[1]
        assert We.shape == (E, L) and Wd.shape == (L, E)
        out[i, :E * L] = np.asarray(We, np.float32).ravel()
        out[i, E * L:2 * E * L] = np.asarray(Wd, np.float32).ravel()
    return out


def unpack_weights(row: np.ndarray, E: int, L: int):
    return row[:E * L].reshape(E, L).copy(), row[E * L:2 * E * L].reshape(L, E).copy()


_WINDOW_TABLES: dict = {}


def window_index(kind: str, T: int, window: int):
    """Bit-exact window tables from the C ABI: returns (starts, lens) int32 (cached per argument set)."""
    hit = _WINDOW_TABLES.get((kind, T, window))
    if hit is not None:
        return hit[0].copy(), hit[1].copy()
    starts, lens = _window_index(kind, T, window)
    _WINDOW_TABLES[(kind, T, window)] = (starts.copy(), lens.copy())
    return starts, lens


def _window_index(kind: str, T: int, window: int):
    k = {"rolling": 0, "expanding": 1, "expost": 2}[kind]
    n = check(lib.hfrep_window_index(k, T, window, None, None, 0))
    starts, lens = np.zeros(n, np.int32), np.zeros(n, np.int32)
    check(lib.hfrep_window_index(k, T, window, ptr(starts), ptr(lens), n))
    return starts, lens


# ------------------------------------------------------------------------------------ device helpers
def _dev(device=None):
    if not torch.cuda.is_available():
        raise _lib.HfrepError("no CUDA device: this package has no CPU fallback")
    return torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")


class _Staging:
    """Pinned host staging buffers, reused: `tensor.pin_memory()` allocates page-locked memory on every call (hundreds
    of microseconds each — most of the host overhead of a small end-to-end step).  A buffer is handed out again only
    after the copy that last read it has completed (one CUDA event per buffer)."""

    def __init__(self):
        self.free = []                       # [(uint8 pinned tensor, event of its last copy)]

    def take(self, nbytes: int):
        for i, (buf, ev) in enumerate(self.free):
            if buf.numel() >= nbytes and ev.query():
                return self.free.pop(i)
        if len(self.free) > 32:
            self.free = [(b, e) for b, e in self.free if not e.query()]      # drop idle buffers; busy ones stay alive
        return torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8).pin_memory(), torch.cuda.Event()

    def give(self, buf, ev):
        self.free.append((buf, ev))


_STAGING = _Staging()


def to_device(a: np.ndarray, device, dtype=None) -> torch.Tensor:
    """Host array -> pinned staging (pooled) -> device, asynchronous on the current stream."""
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    if t.numel() == 0:
        return t.to(device)
    nbytes = t.numel() * t.element_size()
    buf, ev = _STAGING.take(nbytes)
    host = buf[:nbytes].view(t.dtype).reshape(t.shape)
    host.copy_(t)
    out = host.to(device, non_blocking=True)
    ev.record(torch.cuda.current_stream(out.device))
    _STAGING.give(buf, ev)
    return out


_PAD: dict = {}


def to_host(tensors) -> list:
    """Device tensors -> host numpy arrays with ONE device-to-host copy: the tensors' bytes are concatenated on the
    device (one kernel), copied into one page-locked buffer (torch's caching host allocator: no cudaHostAlloc after
    the first call) and handed out as views of it (None entries stay None)."""
    live = [t for t in tensors if t is not None]
    if not live:
        return [None for _ in tensors]
    dev = live[0].device
    zeros = _PAD.get(dev)
    if zeros is None:
        zeros = _PAD[dev] = torch.zeros(16, dtype=torch.uint8, device=dev)
    flat, off = [], 0
    for t in live:                                     # every field starts 16-byte aligned in the packed buffer
        f = t.contiguous().reshape(-1).view(torch.uint8)
        pad = (-off) & 15
        if pad:
            flat.append(zeros[:pad])
            off += pad
        flat.append(f)
        off += f.numel()
    packed = torch.cat(flat)
    host = torch.empty(packed.numel(), dtype=torch.uint8, pin_memory=True)
    host.copy_(packed, non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()
    out, off, it = [], 0, iter(live)
    for t in tensors:
        if t is None:
            out.append(None)
            continue
        tt = next(it)
        off += (-off) & 15
        n = tt.numel() * tt.element_size()
        out.append(host[off:off + n].view(tt.dtype).reshape(tt.shape).numpy())
        off += n
    return out


_SCHED_CACHE: dict = {}


def nadam_schedule(cfg: _lib.FitCfg, n_steps: int, device) -> torch.Tensor:
    key = (cfg.lr, cfg.beta_1, cfg.beta_2, cfg.epsilon, cfg.schedule_decay, cfg.decay_base, n_steps, str(device))
    t = _SCHED_CACHE.get(key)
    if t is None:
        host = np.zeros((n_steps, _lib.SCHED_STRIDE), np.float32)
        check(lib.hfrep_nadam_schedule(C.byref(cfg), n_steps, ptr(host)))
        t = torch.from_numpy(host).to(device)
        _SCHED_CACHE[key] = t
    return t


@dataclasses.dataclass
class FitResult:
    W: torch.Tensor            # f32 [n][w_stride]
    loss: torch.Tensor         # f32 [n][max_epochs]
    val_loss: torch.Tensor
    epochs_run: torch.Tensor   # i32 [n]
    status: torch.Tensor
    grads: torch.Tensor | None
    w_stride: int


_SIDE_STREAMS: dict = {}
_INT_CACHE: dict = {}


def _device_ints(a: np.ndarray, device) -> torch.Tensor:
    """Small int32 task tables live on the device once per distinct content (sweeps reuse them every step)."""
    key = (str(device), a.tobytes())
    t = _INT_CACHE.get(key)
    if t is None:
        if len(_INT_CACHE) > 64:
            _INT_CACHE.clear()
        t = torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(device)
        _INT_CACHE[key] = t
    return t


def _side_streams(device, n):
    pool = _SIDE_STREAMS.setdefault(str(device), [])
    while len(pool) < n:
        pool.append(torch.cuda.Stream(device=device))
    return pool[:n]


def warps_per_task_for(n_tasks: int, device) -> int:
    """Kernel (b) mapping for a job of n_tasks fits running at the same time (all width groups together): the same
    rule hfrep_fit_batched applies to a single call (csrc/fit_kernel.cu, dispatch_mapping)."""
    sms = torch.cuda.get_device_properties(device).multi_processor_count
    if n_tasks * 6 <= sms:
        return _lib.WPT_CLUSTER         # so few tasks that each gets a cluster of six SMs
    return 6 if n_tasks <= 5 * sms else 2


def fit_batched(X: torch.Tensor, task_L, W0: torch.Tensor, *, E: int, T: int, n_fit: int | None = None,
                perm: torch.Tensor | None = None, task_x: torch.Tensor | None = None, dump_step: int | None = None,
                warps_per_task: int = 0, bucket: bool = True, W_out: torch.Tensor | None = None, after_bucket=None,
                shuffle_seeds=None, hist: torch.Tensor | None = None, ints: torch.Tensor | None = None,
                **cfg_kw) -> FitResult:
    """Kernel (b) for a batch of tasks.  X f32 [n_x][T][E] (device, scaled); task_L sequence of ints;
    W0 f32 [n][w_stride] (device).  Tasks are grouped by padded latent width (multiples of 4) and each
    group is launched on its own stream, so a sweep over L runs concurrently."""
    device = X.device
    task_L = np.asarray(task_L, np.int32)
    n = len(task_L)
    w_stride = W0.shape[1]
    n_x = X.shape[0] if X.dim() == 3 else 1
    if n_fit is None:
        n_fit = int(np.floor(T * .75))                       # validation_split=.25, Autoencoder_encapsulate.py:89
    cfg = _lib.fit_cfg(E=E, T=T, n_fit=n_fit, n_x=n_x, warps_per_task=warps_per_task,
                       dump_step=(dump_step or 0), **cfg_kw)
    if n == 0:
        z = lambda *shape, dt=torch.float32: torch.zeros(shape, dtype=dt, device=device)
        return FitResult(z(0, w_stride), z(0, cfg.max_epochs), z(0, cfg.max_epochs), z(0, dt=torch.int32),
                         z(0, dt=torch.int32), None, w_stride)
    steps = cfg.max_epochs * ((n_fit + cfg.batch - 1) // cfg.batch)
    sched = nadam_schedule(cfg, steps, device)
    # order tasks by padded width, widest (slowest) first
    order = np.argsort(-((task_L + 3) // 4), kind="stable") if bucket else np.arange(n)
    inv = np.empty(n, np.int64)
    inv[order] = np.arange(n)
    L_sorted = task_L[order]
    identity = bool(np.array_equal(order, np.arange(n)))      # already widest-first: no gather / scatter needed
    tL = _device_ints(L_sorted, device)
    per_task_perm = perm is not None and perm.dim() == 3
    if identity:
        W0s, tx, perms = W0.contiguous(), task_x, perm
    else:
        order_t = torch.from_numpy(order).to(device)
        W0s = W0.index_select(0, order_t).contiguous()
        tx = task_x.index_select(0, order_t).contiguous() if task_x is not None else None
        perms = perm.index_select(0, order_t).contiguous() if per_task_perm else perm
    perm_stride = perms.shape[1] * perms.shape[2] if per_task_perm else 0
    tseed = None
    if shuffle_seeds is not None:                      # device-side epoch shuffles (Keras shuffle=True), one stream per task
        tseed = _device_ints(np.asarray(shuffle_seeds, np.uint32)[order].view(np.int32), device)

    if W_out is None:
        W_out = torch.empty_like(W0s)      # the kernel writes every float of a row (the tail past the task's own L: zero)
    if after_bucket is not None and not identity:
        raise ValueError("after_bucket needs the tasks ordered widest-first (see run_mode_r)")
    # one allocation for both histories, one for the two int vectors; nothing to initialise (the kernel writes NaN past
    # the last epoch run, and epochs_run / status of every task)
    if hist is None:                                   # (callers that pack results themselves pass both: SweepPlan)
        hist = torch.empty((2, n, cfg.max_epochs), dtype=torch.float32, device=device)
    loss, val = hist[0], hist[1]
    if ints is None:
        ints = torch.empty((2, n), dtype=torch.int32, device=device)
    epochs_run, status = ints[0], ints[1]
    grads = torch.zeros_like(W0s) if dump_step is not None else None

    LP = (L_sorted + 3) // 4 * 4
    bounds = [0] + [i for i in range(1, n) if LP[i] != LP[i - 1]] + [n] if bucket else [0, n]
    groups = [(bounds[i], bounds[i + 1]) for i in range(len(bounds) - 1)]
    job_wpt = 0
    if warps_per_task == 0 and len(groups) > 1:
        job_wpt = cfg.warps_per_task = warps_per_task_for(n, device)   # the groups run concurrently: map by the whole job's size
        sms = torch.cuda.get_device_properties(device).multi_processor_count
    main = torch.cuda.current_stream(device)
    streams = _side_streams(device, len(groups)) if len(groups) > 1 else [main]
    ready = torch.cuda.Event()
    ready.record(main)
    for (a, b), s in zip(groups, streams):
        if s is not main:
            s.wait_event(ready)
        with torch.cuda.stream(s):
            cfg.L_max = int(L_sorted[a:b].max())
            if job_wpt == 2:                    # narrow widths of a very large job: one warp per task (measured: L <= 8 at
                cfg.warps_per_task = 1 if (cfg.L_max <= 8 and n >= 24 * sms) else 2   # 4096 tasks 4.8 / 6.5 vs 6.4 / 7.6 ms)
            off = lambda t, a=a: None if t is None else C.c_void_p(t.data_ptr() + a * t.stride(0) * t.element_size())
            check(lib.hfrep_fit_batched(
                C.byref(cfg), ptr(X), b - a, off(tL), off(tx), off(W0s), w_stride,
                off(perms) if per_task_perm else ptr(perms), perm_stride, off(tseed), ptr(sched), off(W_out), off(loss),
                off(val), off(epochs_run), off(status), off(grads), C.c_void_p(s.cuda_stream)))
            if after_bucket is not None:
                after_bucket(a, b, s.cuda_stream)      # e.g. this bucket's evaluation, right behind its fit
        if s is not main:
            done = torch.cuda.Event()
            done.record(s)
            main.wait_event(done)
    if identity:
        sel = lambda t: t
    else:
        inv_t = torch.from_numpy(inv).to(device)
        sel = lambda t: None if t is None else t.index_select(0, inv_t)
    return FitResult(sel(W_out), sel(loss), sel(val), sel(epochs_run), sel(status), sel(grads), w_stride)


@dataclasses.dataclass
class EvalResult:
    ante: torch.Tensor         # f64 [n][n_oos][F]
    post: torch.Tensor | None  # None when no sigma was given
    turnover: torch.Tensor     # f64 [n][F]
    metrics: torch.Tensor      # f64 [n][2][N_STATS][F]   _lib.STAT_NAMES of ante, then of post
    quality: torch.Tensor      # f64 [n][2+2*n_pre]
    mask: torch.Tensor         # u8  [n][n_oos][E]
    beta: torch.Tensor         # f64 [n][MAX_L][F]
    norm: torch.Tensor         # f64 [n][F]
    status: torch.Tensor
    weights: torch.Tensor | None


def rolling_sigma(R64: torch.Tensor, starts: torch.Tensor, window: int) -> torch.Tensor:
    T, E = R64.shape
    out = torch.empty((len(starts), E), dtype=torch.float64, device=R64.device)
    check(lib.hfrep_rolling_sigma(ptr(R64), T, E, ptr(starts), len(starts), window, ptr(out), stream_ptr()))
    return out


def predict_metrics(Xtr64, Xte64, Yte64, rf64, sigma, task_L, W, *, window=24, want_oos=True,
                    want_weights=False, alpha=.2, cost_param=.05, phi=.5, task_row0=None, T_test=None,
                    want_quality=True, launch=True, beta_rolling=False, out: dict | None = None):
    """Kernel (c) for a batch of trained tasks (all tensors on the device, float64 unless noted).
    launch=False only allocates the result and returns (result, launcher): launcher(a, b, stream) scores tasks
    [a, b) on `stream` — used to start a width bucket's evaluation right behind its fit, concurrently with the
    other buckets' fits."""
    device = Xte64.device
    n = len(task_L)
    E = Xte64.shape[1]
    T_test = Xte64.shape[0] if T_test is None else T_test      # rolling refit: rows per task, not the history
    F = Yte64.shape[1]
    cfg = _lib.eval_cfg(E=E, F=F, T_train=(Xtr64.shape[0] if Xtr64 is not None else 0), T_test=T_test, window=window,
                        want_oos=int(want_oos), alpha=alpha, cost_param=cost_param, phi=phi, beta_rolling=int(beta_rolling))
    want_weights = want_weights or beta_rolling       # rolling betas keep the per-month weights in that buffer
    n_oos, n_pre = T_test - window - 1, T_test - 2
    tL = task_L if isinstance(task_L, torch.Tensor) else _device_ints(np.asarray(task_L, np.int32), device)
    f64 = dict(dtype=torch.float64, device=device)
    o = out or {}                                      # preallocated result fields (SweepPlan: views of its download buffer)
    nan_ = lambda key, shape: o[key].fill_(float("nan")) if key in o else torch.full(shape, float("nan"), **f64)
    r = EvalResult(
        ante=o.get("ante", None) if "ante" in o else torch.empty((n, n_oos, F), **f64),
        post=(o["post"] if "post" in o else torch.empty((n, n_oos, F), **f64)) if sigma is not None else None,
        turnover=o["turnover"] if "turnover" in o else torch.empty((n, F), **f64), metrics=nan_("metrics", (n, 2, _lib.N_STATS, F)),
        quality=nan_("quality", (n, 2 + 2 * n_pre)) if want_quality else None,
        mask=torch.empty((n, n_oos, E), dtype=torch.uint8, device=device),
        beta=torch.empty((n, _lib.MAX_L, F), **f64), norm=torch.empty((n, F), **f64),
        status=o["eval_status"].zero_() if "eval_status" in o else torch.zeros(n, dtype=torch.int32, device=device),
        weights=torch.empty((n, n_oos, E, F), **f64) if want_weights else None)

    def launcher(a, b, stream):
        off = lambda t: None if t is None else C.c_void_p(t.data_ptr() + a * t.stride(0) * t.element_size())
        check(lib.hfrep_predict_metrics(
            C.byref(cfg), ptr(Xtr64), ptr(Xte64), ptr(Yte64), ptr(rf64), ptr(sigma), b - a, off(tL), off(task_row0), off(W),
            W.shape[1], off(r.ante), off(r.post), off(r.turnover), off(r.metrics), off(r.quality), off(r.mask),
            off(r.weights), off(r.beta), off(r.norm), off(r.status), C.c_void_p(stream)))

    if not launch:
        return r, launcher
    launcher(0, n, torch.cuda.current_stream(device).cuda_stream)
    return r


def window_gather(R: torch.Tensor, starts, length: int, col_lo: int = 0, ncols: int | None = None,
                  minmax: bool = False) -> torch.Tensor:
    """Kernel (a): out[w] = R[starts[w]:starts[w]+length, col_lo:col_lo+ncols] (f32, device).  R float32 or float64
    (device); a float64 source is scaled in float64 and rounded once, as MinMaxScaler + Keras' cast do."""
    T, Cc = R.shape
    ncols = Cc - col_lo if ncols is None else ncols
    st = starts if isinstance(starts, torch.Tensor) else torch.from_numpy(np.asarray(starts, np.int32)).to(R.device)
    if len(st) and (int(st.min()) < 0 or int(st.max()) + length > T):
        raise _lib.HfrepError("window_gather: a window reaches outside the returns matrix")
    out = torch.empty((len(st), length, ncols), dtype=torch.float32, device=R.device)
    if R.dtype not in (torch.float32, torch.float64) or not R.is_contiguous():
        raise _lib.HfrepError("window_gather: R must be a contiguous float32 or float64 matrix")
    fn = lib.hfrep_window_gather if R.dtype == torch.float32 else lib.hfrep_window_gather_f64
    check(fn(ptr(R), T, Cc, ptr(st), len(st), length, col_lo, ncols, int(minmax), ptr(out), stream_ptr()))
    return out


# ------------------------------------------------------------------------------------ Mode R front-end
class ReplicationData:
    """The reference's data split, resident on one GPU: train/test halves of the ETF and fund
    matrices (autoencoder_v4.ipynb:112), the MinMax-scaled train block (Autoencoder_encapsulate.py:65),
    the raw test block (:67), rf aligned with the test block and the rolling ex-post volatilities of
    the last n_oos+window ETF rows (helper.py:121; task-independent, computed once)."""

    def __init__(self, etf, hfd, rf, window: int = 24, device=None, test_size: float = .5):
        etf, hfd = np.asarray(etf, np.float64), np.asarray(hfd, np.float64)
        rf = np.asarray(rf, np.float64).reshape(-1)
        assert len(etf) == len(hfd) == len(rf)
        self._shape(len(etf), etf.shape[1], hfd.shape[1], window, _dev(device), test_size)
        with torch.cuda.device(self.device):
            # one host-to-device copy for the whole data set: [scaled train block | test ETFs | test funds | test rf]
            self._bind(to_device(self.pack_host(etf, hfd, rf, self.n_train), self.device))

    def _shape(self, T, E, F, window, device, test_size):
        self.device = device
        self.n_train, self.n_test = split_rows(T, test_size)
        self.E, self.F, self.window = E, F, window
        self.n_oos = self.n_test - window - 1
        self.n_image = self.n_train * E + self.n_test * (E + F + 1)      # float64 values of the device image
        self.h2d_bytes = 8 * self.n_image

    @staticmethod
    def pack_host(etf, hfd, rf, ntr, out=None):
        """The float64 host image of a data set: MinMax-scaled train ETFs, then the test ETFs, funds and rf."""
        return np.concatenate([minmax_scale(etf[:ntr]).ravel(), etf[ntr:].ravel(), hfd[ntr:].ravel(), rf[ntr:]], out=out)

    @classmethod
    def from_packed(cls, d: torch.Tensor, T: int, E: int, F: int, window: int = 24, test_size: float = .5):
        """The same object over an image (`pack_host`) already on the device — no host work, no synchronisation."""
        self = cls.__new__(cls)
        self._shape(T, E, F, window, d.device, test_size)
        with torch.cuda.device(self.device):
            self._bind(d)
        return self

    def _bind(self, d):
        ntr, nte, E, F, window = self.n_train, self.n_test, self.E, self.F, self.window
        o1, o2, o3 = ntr * E, ntr * E + nte * E, ntr * E + nte * E + nte * F
        self.Xtr64 = d[:o1].view(ntr, E)
        self.Xtr32 = self.Xtr64.to(torch.float32).reshape(1, ntr, E).contiguous()
        self.Xte64 = d[o1:o2].view(nte, E)
        self.Yte64 = d[o2:o3].view(nte, F)
        self.rf64 = d[o3:]
        tail = self.Xte64[nte - self.n_oos - window:]        # the last n_oos+window ETF rows (Autoencoder_encapsulate.py:206)
        starts, _ = window_index("expost", self.n_oos + window, window)               # helper.py:120
        self.sigma = rolling_sigma(tail, _device_ints(starts, self.device), window)


def fit_and_score(data: "ReplicationData", task_L, W0d: torch.Tensor, *, perm=None, want_oos=True, want_weights=False,
                  warps_per_task=0, out: dict | None = None, **fit_kw):
    """Kernel (b) then kernel (c) for tasks ordered widest-first: every width bucket's evaluation is queued on the
    bucket's own stream right behind its fit, so it overlaps the fits of the other buckets."""
    o = out or {}
    W_out = o["W"] if "W" in o else torch.empty_like(W0d)
    ev, score = predict_metrics(data.Xtr64, data.Xte64, data.Yte64, data.rf64, data.sigma, task_L, W_out,
                                window=data.window, want_oos=want_oos, want_weights=want_weights,
                                alpha=fit_kw.get("alpha", .2), launch=False, out=out)
    fit = fit_batched(data.Xtr32, task_L, W0d, E=data.E, T=data.n_train, perm=perm, warps_per_task=warps_per_task,
                      W_out=W_out, after_bucket=score, hist=o.get("hist"), ints=o.get("ints"), **fit_kw)      # fit_kw may carry shuffle_seeds
    return fit, ev


def mode_r_tasks(latent_dims, seeds):
    """The (L, seed) grid of a Mode-R job in its canonical order: widest (slowest) tasks first — the launch order of the
    width buckets, and the order the cyclic multi-GPU split deals from so that every rank gets the same mix of widths."""
    return sorted(((int(L), int(s)) for s in seeds for L in latent_dims), key=lambda t: -((t[0] + 3) // 4))


def run_mode_r(data: ReplicationData, latent_dims, seeds=(123,), *, W0=None, perm=None, shuffle=False, want_oos=True,
               want_weights=False, warps_per_task=0, sharded=False, timings: dict | None = None, **fit_kw):
    """The notebook's sweep (autoencoder_v4.ipynb:158-165, :896-901) as one batched GPU pass:
    for every (L, seed): fit, IS/OOS quality, ante, turnover, post, metrics.  shuffle=False is the deterministic row
    order used for parity; shuffle=True reshuffles the fit rows every epoch on the device (Keras' default).
    Returns (tasks, FitResult, EvalResult); tasks[i] = (L, seed).

    sharded=True inside an initialised torch.distributed job (one process per GPU, `data` resident on this rank's GPU):
    the task list is dealt cyclically over the ranks (distributed.shard), every rank fits and scores its own tasks — no
    collective on the data path — and ONE all-gather of packed result records (distributed.gather_records, NCCL over
    NVLink) leaves the complete results, in global task order, on every rank.  W0 / per-task perm, when given, cover
    the whole grid in `mode_r_tasks` order.  `timings` receives compute_ms / gather_ms (CUDA events) when given."""
    from . import distributed
    tasks = mode_r_tasks(latent_dims, seeds)
    rank, world = distributed.world() if sharded else (0, 1)
    mine = distributed.shard(len(tasks), rank, world)
    local = [tasks[i] for i in mine]
    E = data.E
    L_max = max(L for L, _ in tasks)
    w_stride = 2 * E * L_max
    with torch.cuda.device(data.device):
        ev_t = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if timings is not None else None
        if W0 is None:
            W0 = pack_weights([initial_weights(s, E, L) for L, s in local], E, w_stride)
        elif world > 1:
            W0 = np.asarray(W0, np.float32)[mine]
        W0d = to_device(np.asarray(W0, np.float32), data.device)
        if perm is not None and world > 1 and np.asarray(perm).ndim == 3:
            perm = np.asarray(perm)[mine]
        permd = None if perm is None else to_device(np.asarray(perm, np.int32), data.device)
        task_L = [L for L, _ in local]
        if shuffle and perm is None:        # Keras' default shuffle=True: one device-side stream per (L, seed)
            fit_kw["shuffle_seeds"] = [(s * 1000003 + L) & 0xFFFFFFFF for L, s in local]
        if ev_t:
            ev_t[0].record()
        fit, ev = fit_and_score(data, task_L, W0d, perm=permd, want_oos=want_oos, want_weights=want_weights,
                                warps_per_task=warps_per_task, **fit_kw)
        if ev_t:
            ev_t[1].record()
        if world > 1:
            fields = [fit.W, fit.loss, fit.val_loss, fit.epochs_run, fit.status, fit.grads, ev.ante, ev.post, ev.turnover,
                      ev.metrics, ev.quality, ev.mask, ev.beta, ev.norm, ev.status, ev.weights]
            g = distributed.gather_records(fields, len(tasks), rank, world)
            fit = FitResult(g[0], g[1], g[2], g[3], g[4], g[5], fit.w_stride)
            ev = EvalResult(ante=g[6], post=g[7], turnover=g[8], metrics=g[9], quality=g[10], mask=g[11], beta=g[12],
                            norm=g[13], status=g[14], weights=g[15])
        if ev_t:
            ev_t[2].record()
            ev_t[2].synchronize()
            timings["compute_ms"] = ev_t[0].elapsed_time(ev_t[1])
            timings["gather_ms"] = ev_t[1].elapsed_time(ev_t[2])
            timings["record_bytes_per_task"] = int(sum(t[0].numel() * t.element_size() for t in (fit.W, fit.loss, fit.val_loss, fit.epochs_run, fit.status, ev.ante, ev.post, ev.turnover, ev.metrics, ev.quality, ev.mask, ev.beta, ev.norm, ev.status) if t is not None))
    return tasks, fit, ev


class SweepResult:
    """Host views of one `SweepPlan.run` (page-locked memory owned by the plan: valid until its next run)."""
    __slots__ = ("tasks", "ante", "post", "turnover", "metrics", "quality", "eval_status", "epochs_run", "status", "loss",
                 "val_loss", "W", "h2d_bytes", "d2h_bytes")


class SweepPlan:
    """A Mode-R sweep of FIXED shape (history length, ETF and fund counts, widths, seeds) recorded once as a CUDA graph:
    one host-to-device copy of the data image and the initial kernels, the scaled block's float32 copy, the rolling
    volatilities, every width bucket's fit and evaluation on its own captured branch, the packing of the results and one
    device-to-host copy.  `run(etf, hfd, rf)` then costs one host memcpy into page-locked memory, one graph launch and one
    wait — the per-call launches, event records and result allocations of `run_mode_r` (≈ 0.6 ms of host time on a 5.5 ms
    sweep) are gone.  Results are bit-identical to `run_mode_r` on the same inputs (tests/test_gpu_parity.py)."""

    def __init__(self, T: int, E: int, F: int, latent_dims, seeds=(123,), *, window: int = 24, device=None,
                 test_size: float = .5, shuffle: bool = False, want_oos: bool = True, **fit_kw):
        self.device = dev = _dev(device)
        self.T, self.E, self.F, self.window, self.test_size = T, E, F, window, test_size
        self.tasks = mode_r_tasks(latent_dims, seeds)
        task_L = [L for L, _ in self.tasks]
        n = len(self.tasks)
        self.w_stride = 2 * E * max(task_L)
        self.n_train, n_test = split_rows(T, test_size)
        n_img = self.n_train * E + n_test * (E + F + 1)
        if shuffle:
            fit_kw["shuffle_seeds"] = [(s * 1000003 + L) & 0xFFFFFFFF for L, s in self.tasks]
        with torch.cuda.device(dev):
            # page-locked input image: [data image (f64) | initial kernels (f32)], and its device twin
            self._in_host = torch.empty(8 * n_img + 4 * n * self.w_stride, dtype=torch.uint8, pin_memory=True)
            self._img_host = self._in_host[:8 * n_img].view(torch.float64).numpy()
            self._w0_host = self._in_host[8 * n_img:].view(torch.float32).reshape(n, self.w_stride).numpy()
            self._w0_host[:] = pack_weights([initial_weights(s, E, L) for L, s in self.tasks], E, self.w_stride)
            self._default_w0 = self._w0_host.copy()
            self._img_host[:] = 0.
            in_dev = torch.empty_like(self._in_host, device=dev)

            # the results a caller downloads are views of ONE device buffer, so the kernels write the packed image themselves
            n_test = split_rows(T, test_size)[1]
            n_oos, n_pre = n_test - window - 1, n_test - 2
            epochs = _lib.fit_cfg(E=E, T=self.n_train, n_fit=1, **{k: v for k, v in fit_kw.items() if k != "shuffle_seeds"}).max_epochs
            fields = [("ante", torch.float64, (n, n_oos, F)), ("post", torch.float64, (n, n_oos, F)), ("turnover", torch.float64, (n, F)),
                      ("metrics", torch.float64, (n, 2, _lib.N_STATS, F)), ("quality", torch.float64, (n, 2 + 2 * n_pre)),
                      ("eval_status", torch.int32, (n,)), ("ints", torch.int32, (2, n)), ("hist", torch.float32, (2, n, epochs)),
                      ("W", torch.float32, (n, self.w_stride))]
            carve, off = {}, 0
            for name, dt, shape in fields:
                off += (-off) & 15
                carve[name] = (off, dt, shape)
                off += int(np.prod(shape)) * torch.empty((), dtype=dt).element_size()
            out_dev = torch.zeros(off, dtype=torch.uint8, device=dev)
            views = {k: out_dev[o:o + int(np.prod(sh)) * torch.empty((), dtype=dt).element_size()].view(dt).reshape(sh)
                     for k, (o, dt, sh) in carve.items()}
            self._out_host = torch.empty(off, dtype=torch.uint8, pin_memory=True)
            eb = 4 * n * epochs                                  # bytes of one history
            o_i, o_h = carve["ints"][0], carve["hist"][0]
            self._layout = [carve["ante"], carve["post"], carve["turnover"], carve["metrics"], carve["quality"], carve["eval_status"],
                            (o_i, torch.int32, (n,)), (o_i + 4 * n, torch.int32, (n,)),
                            (o_h, torch.float32, (n, epochs)), (o_h + eb, torch.float32, (n, epochs)), carve["W"]]

            def body():
                in_dev.copy_(self._in_host, non_blocking=True)
                data = ReplicationData.from_packed(in_dev[:8 * n_img].view(torch.float64), T, E, F, window, test_size)
                W0d = in_dev[8 * n_img:].view(torch.float32).reshape(n, self.w_stride)
                fit, ev = fit_and_score(data, task_L, W0d, want_oos=want_oos, out=views, **fit_kw)
                assert fit.W.data_ptr() == views["W"].data_ptr() and fit.loss.data_ptr() == views["hist"].data_ptr() \
                    and ev.ante.data_ptr() == views["ante"].data_ptr(), "results not written in place"
                self._out_host.copy_(out_dev, non_blocking=True)

            # once eagerly on a well-formed image: fills the small caches (schedules, index tables), loads the kernels
            rng = np.random.default_rng(0)
            self._img_host[:] = rng.uniform(.01, .02, n_img)
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                body()
            side.synchronize()
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph, stream=side, capture_error_mode="thread_local"):
                body()
            self._stream = side
        self.h2d_bytes, self.d2h_bytes = self._in_host.numel(), off

    def run(self, etf, hfd, rf, W0=None) -> SweepResult:
        etf, hfd = np.asarray(etf, np.float64), np.asarray(hfd, np.float64)
        rf = np.asarray(rf, np.float64).reshape(-1)
        if etf.shape != (self.T, self.E) or hfd.shape != (self.T, self.F) or rf.shape != (self.T,):
            raise _lib.HfrepError(f"SweepPlan.run: this plan was recorded for etf {(self.T, self.E)}, hfd {(self.T, self.F)}")
        ReplicationData.pack_host(etf, hfd, rf, self.n_train, out=self._img_host)
        self._w0_host[:] = self._default_w0 if W0 is None else np.asarray(W0, np.float32)
        with torch.cuda.stream(self._stream):
            self._graph.replay()
        self._stream.synchronize()
        r = SweepResult()
        r.tasks, r.h2d_bytes, r.d2h_bytes = self.tasks, self.h2d_bytes, self.d2h_bytes
        names = ("ante", "post", "turnover", "metrics", "quality", "eval_status", "epochs_run", "status", "loss", "val_loss", "W")
        for name, (o, dt, shape) in zip(names, self._layout):
            nb = int(np.prod(shape)) * torch.empty((), dtype=dt).element_size()
            setattr(r, name, self._out_host[o:o + nb].view(dt).reshape(shape).numpy())
        return r


# ------------------------------------------------------------------------------------ Mode W front-end
def fit_windows(R32: torch.Tensor, win_starts, T_win: int, task_L, task_win, W0: torch.Tensor, *, E: int, col_lo: int = 0,
                perm: torch.Tensor | None = None, warps_per_task: int = 0, shuffle_seeds=None, **cfg_kw):
    """hfrep_fit_windows: kernel (a) gathers + MinMax-scales every training window into a scratch buffer, kernel (b)
    fits task i on window task_win[i].  Returns (FitResult, scratch blocks f32 [n_windows][T_win][E])."""
    device = R32.device
    T_all, Cc = R32.shape
    task_L = np.asarray(task_L, np.int32)
    n = len(task_L)
    n_fit = int(np.floor(T_win * .75))
    cfg = _lib.fit_cfg(E=E, T=T_win, n_fit=n_fit, L_max=int(task_L.max()), warps_per_task=warps_per_task, **cfg_kw)
    steps = cfg.max_epochs * ((n_fit + cfg.batch - 1) // cfg.batch)
    sched = nadam_schedule(cfg, steps, device)
    starts = _device_ints(np.asarray(win_starts, np.int32), device)
    n_win = len(starts)
    ws_bytes = lib.hfrep_fit_windows_workspace(n_win, T_win, E)
    ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=device)
    tL, tw = _device_ints(task_L, device), _device_ints(np.asarray(task_win, np.int32), device)
    W_out = torch.zeros_like(W0)
    loss = torch.full((n, cfg.max_epochs), float("nan"), dtype=torch.float32, device=device)
    val = torch.full((n, cfg.max_epochs), float("nan"), dtype=torch.float32, device=device)
    epochs_run = torch.zeros(n, dtype=torch.int32, device=device)
    status = torch.zeros(n, dtype=torch.int32, device=device)
    per_task_perm = perm is not None and perm.dim() == 3
    tseed = None if shuffle_seeds is None else _device_ints(np.asarray(shuffle_seeds, np.uint32).view(np.int32), device)
    check(lib.hfrep_fit_windows(C.byref(cfg), ptr(R32), T_all, Cc, col_lo, ptr(starts), n_win, n, ptr(tL), ptr(tw),
                                ptr(W0), W0.shape[1], ptr(perm), perm.shape[1] * perm.shape[2] if per_task_perm else 0,
                                ptr(tseed), ptr(sched), ptr(ws), ws_bytes, ptr(W_out), ptr(loss), ptr(val), ptr(epochs_run),
                                ptr(status), stream_ptr()))
    return FitResult(W_out, loss, val, epochs_run, status, None, W0.shape[1]), ws.view(n_win, T_win, E)


def run_mode_w_wide(etf, hfd, rf, *, train_window: int, latent_dim: int, seeds=(123,), ols_window: int = 24, device=None,
                    W0=None, windows=None, timings: dict | None = None, **fit_kw):
    """Rolling refit for wide shapes (hundreds of ETFs, thousands of funds: BASELINE configs[3]).  Same task definition
    and alignment as run_mode_w; kernel (a) builds the MinMax-scaled training windows, hfrep_fit_wide trains on them with
    weights in a global workspace, hfrep_predict_wide regresses every fund on the latent factors and returns the replica
    return of row t+1.  `windows`: optional subset of window indices (a rank's share of the job, a bounded sample); only the
    rows of the history those windows touch are uploaded, so a rank that takes a run of consecutive windows
    (distributed.shard_blocks) uploads its own stretch of the histories and nothing else."""
    dev = _dev(device)
    src_f32 = np.asarray(etf).dtype == np.float32       # float32 history: windows are cut from it as it is; otherwise the
    etf = np.asarray(etf, np.float64)                   # MinMax runs on the float64 values (as the reference's AE does)
    hfd = np.asarray(hfd, np.float64)
    rf = np.asarray(rf, np.float64).reshape(-1)
    T, E = etf.shape
    F = hfd.shape[1]
    L = latent_dim
    ts = np.arange(train_window, T - 1)
    if windows is not None:
        ts = ts[np.asarray(windows)]
    n_win = len(ts)
    tasks = [(wi, int(s)) for s in seeds for wi in range(n_win)]
    n = len(tasks)
    t_out = ts
    if n_win:                                            # only the stretch of the history these windows touch travels to the GPU:
        r_lo, r_hi = int(ts.min()) - train_window, int(ts.max()) + 2      # rows [t - train_window, t + 2) of every window
        etf, hfd, rf, ts, T = etf[r_lo:r_hi], hfd[r_lo:r_hi], rf[r_lo:r_hi], ts - r_lo, r_hi - r_lo
    with torch.cuda.device(dev):
        etf_d = to_device(etf, dev)
        Rsrc = etf_d.to(torch.float32) if src_f32 else etf_d
        blocks = window_gather(Rsrc, (ts - train_window).astype(np.int32), train_window, 0, E, minmax=True)
        if W0 is None:
            W0 = pack_weights([initial_weights(s * 100003 + wi, E, L) for wi, s in tasks], E, 2 * E * L)
        W0d = to_device(np.asarray(W0, np.float32), dev)
        n_fit = int(np.floor(train_window * .75))
        cfg = _lib.fit_cfg(E=E, T=train_window, n_fit=n_fit, L_max=min(L, _lib.MAX_L), n_x=n_win, **fit_kw)
        steps = cfg.max_epochs * ((n_fit + cfg.batch - 1) // cfg.batch)
        sched = nadam_schedule(cfg, steps, dev)
        ws_bytes = lib.hfrep_fit_wide_workspace(n, E, L)
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=dev)
        W_out = torch.empty_like(W0d)
        loss = torch.full((n, cfg.max_epochs), float("nan"), dtype=torch.float32, device=dev)
        val = torch.full((n, cfg.max_epochs), float("nan"), dtype=torch.float32, device=dev)
        epochs_run = torch.zeros(n, dtype=torch.int32, device=dev)
        status = torch.zeros(n, dtype=torch.int32, device=dev)
        tw = _device_ints(np.asarray([wi for wi, _ in tasks], np.int32), dev)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if timings is not None else None
        if ev:
            ev[0].record()
        check(lib.hfrep_fit_wide(C.byref(cfg), ptr(blocks), n, L, ptr(tw), ptr(W0d), None, 0, ptr(sched), ptr(ws), ws_bytes,
                                 ptr(W_out), ptr(loss), ptr(val), ptr(epochs_run), ptr(status), stream_ptr()))
        if ev:
            ev[1].record()
        fit = FitResult(W_out, loss, val, epochs_run, status, None, 2 * E * L)
        row0 = _device_ints((ts - ols_window)[[wi for wi, _ in tasks]].astype(np.int32), dev)
        hfd_d, rf_d = to_device(hfd, dev), to_device(rf, dev)
        replica = torch.empty((n, F), dtype=torch.float64, device=dev)
        est = torch.zeros(n, dtype=torch.int32, device=dev)
        check(lib.hfrep_predict_wide(ptr(etf_d), ptr(hfd_d), ptr(rf_d), T, E, F, L, ols_window, fit_kw.get("alpha", .2), n,
                                     ptr(row0), ptr(W_out), ptr(replica), ptr(est), stream_ptr()))
        if ev:                                   # kernel times of the fit and of its input copies + evaluation (CUDA events)
            ev[2].record()
            ev[2].synchronize()
            timings["fit_ms"] = ev[0].elapsed_time(ev[1])
            timings["eval_ms"] = ev[1].elapsed_time(ev[2])
    return dict(t=t_out, replica=replica.view(len(seeds), n_win, F), fit=fit, eval_status=est, blocks=blocks, tasks=tasks,
                h2d_bytes=int(etf.nbytes + hfd.nbytes + rf.nbytes))


def run_mode_w(etf, hfd, rf, *, train_window: int, latent_dim: int, seeds=(123,), ols_window: int = 24, device=None,
               W0=None, sharded=False, **fit_kw):
    """Rolling refit (north_star's per-(window, seed) task; SURVEY §8d Mode W — an extension, the reference fits once):
    for t = train_window .. T-2 fit the autoencoder on rows [t-train_window, t) (MinMax on those rows, 75/25 split),
    then follow the reference's own alignment (Autoencoder_encapsulate.py:149-150,164,183): OLS on rows [t-ols_window, t),
    mask from row t, replica for row t+1.  Returns dict(t, replica [seed][window][F], weights, fit).
    sharded=True (inside a torch.distributed job, one process per GPU): the (window, seed) tasks are dealt cyclically
    over the ranks, the returns matrices are replicated, and one all-gather of packed records returns the complete
    replica / weights / kernels on every rank (see run_mode_r)."""
    from . import distributed
    dev = _dev(device)
    src_f32 = np.asarray(etf).dtype == np.float32       # see run_mode_w_wide
    etf = np.asarray(etf, np.float64)
    hfd = np.asarray(hfd, np.float64)
    rf = np.asarray(rf, np.float64).reshape(-1)
    T, E = etf.shape
    assert train_window >= ols_window and T >= train_window + 2
    ts = np.arange(train_window, T - 1)
    n_win = len(ts)
    all_tasks = [(wi, int(s)) for s in seeds for wi in range(n_win)]
    rank, world = distributed.world() if sharded else (0, 1)
    mine = distributed.shard(len(all_tasks), rank, world)
    tasks = [all_tasks[i] for i in mine]
    task_L = [latent_dim] * len(tasks)
    w_stride = 2 * E * latent_dim
    with torch.cuda.device(dev):
        etf_d = to_device(etf, dev)
        if W0 is None:
            W0 = pack_weights([initial_weights(s * 100003 + wi, E, latent_dim) for wi, s in tasks], E, w_stride)
        elif world > 1:
            W0 = np.asarray(W0, np.float32)[mine]
        W0d = to_device(np.asarray(W0, np.float32), dev)
        task_win = [wi for wi, _ in tasks]
        if src_f32:        # one call: kernel (a) cuts + scales the float32 windows into a workspace, kernel (b) fits
            fit, blocks = fit_windows(etf_d.to(torch.float32), ts - train_window, train_window, task_L, task_win, W0d, E=E,
                                      **fit_kw)
        else:              # float64 history: kernel (a)'s float64 path, then kernel (b) on the blocks
            blocks = window_gather(etf_d, (ts - train_window).astype(np.int32), train_window, 0, E, minmax=True)
            fit = fit_batched(blocks, task_L, W0d, E=E, T=train_window, task_x=_device_ints(np.asarray(task_win, np.int32), dev),
                              **fit_kw)
        row0 = _device_ints((ts - ols_window)[[wi for wi, _ in tasks]].astype(np.int32), dev)
        ev = predict_metrics(None, etf_d, to_device(hfd, dev), to_device(rf, dev), None, task_L, fit.W,
                             window=ols_window, want_oos=False, want_quality=False, want_weights=True,
                             task_row0=row0, T_test=ols_window + 2, alpha=fit_kw.get("alpha", .2))
        F = hfd.shape[1]
        if world > 1:
            g = distributed.gather_records([ev.ante, ev.weights, ev.status, fit.W, fit.loss, fit.val_loss, fit.epochs_run,
                                            fit.status], len(all_tasks), rank, world)
            ev = dataclasses.replace(ev, ante=g[0], weights=g[1], status=g[2])
            fit = FitResult(g[3], g[4], g[5], g[6], g[7], None, fit.w_stride)
        replica = ev.ante.view(len(seeds), n_win, F)
        weights = ev.weights.view(len(seeds), n_win, E, F)
    return dict(t=ts, replica=replica, weights=weights, fit=fit, eval=ev, blocks=blocks, tasks=all_tasks)
