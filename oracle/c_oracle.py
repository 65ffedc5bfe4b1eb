"""ctypes wrapper of oracle/c/fit_oracle.c — TEST INFRASTRUCTURE (see oracle/__init__.py): the plain-C float32
restatement of the reference's fit.  `canonical=True` runs the canonical summation order of DESIGN.md §5 (bit-identical
to oracle/fitcore.py's canonical order and to every mapping of the CUDA fit kernel); `canonical=False` the natural
loop order at full optimisation — the "best-effort CPU" baseline (all cores, OpenMP over tasks) of bench.py."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _cpu_key():
    """-march=native code only runs on the CPU it was built for: one library per host CPU model."""
    import hashlib
    try:
        lines = [l for l in open("/proc/cpuinfo") if l.startswith(("model name", "flags"))][:2]
    except OSError:
        lines = []
    return hashlib.md5("".join(lines).encode()).hexdigest()[:10]


LIB = os.path.join(_HERE, "_build", f"libfit_oracle_{_cpu_key()}.so")


class Cfg(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("E", "L", "T", "n_fit", "batch", "epochs", "patience", "early_stop")] + \
               [(n, C.c_float) for n in ("alpha", "lr", "b1", "b2", "eps", "decay", "base")]


def build():
    subprocess.run(["make", "-s", "-C", os.path.join(_HERE, "c"), f"LIBNAME={os.path.basename(LIB)}"], check=True)


def _lib():
    if not os.path.isfile(LIB):
        build()
    return C.CDLL(LIB)


def _cfg(E, L, T, epochs, batch=48, patience=5, early_stop=True, validation_split=.25):
    return Cfg(E, L, T, int(np.floor(T * (1 - validation_split))), batch, epochs, patience, int(early_stop),
               .2, 1e-3, .9, .999, 1e-7, .004, .96)


def fit(x, We0, Wd0, epochs=1000, perm=None, canonical=True, grad_step=None, **kw):
    x = np.ascontiguousarray(x, np.float32)
    T, E = x.shape
    L = We0.shape[1]
    cfg = _cfg(E, L, T, epochs, **kw)
    We, Wd = np.array(We0, np.float32, copy=True), np.array(Wd0, np.float32, copy=True)
    loss, val = np.zeros(epochs, np.float32), np.zeros(epochs, np.float32)
    p = None if perm is None else np.ascontiguousarray(perm, np.int32).ctypes.data_as(C.c_void_p)
    fn = _lib().fit_one_canonical if canonical else _lib().fit_one_fast
    fn.restype = C.c_int
    grad = np.zeros(2 * E * L, np.float32) if grad_step is not None else None
    n = fn(C.byref(cfg), x.ctypes.data_as(C.c_void_p), We.ctypes.data_as(C.c_void_p), Wd.ctypes.data_as(C.c_void_p),
           p, loss.ctypes.data_as(C.c_void_p), val.ctypes.data_as(C.c_void_p),
           None if grad is None else grad.ctypes.data_as(C.c_void_p), C.c_int(-1 if grad_step is None else grad_step))
    out = dict(We=We, Wd=Wd, loss=loss[:n], val_loss=val[:n], epochs_run=n)
    if grad is not None:
        out["grads"] = {grad_step: (grad[:E * L].reshape(E, L).copy(), grad[E * L:].reshape(L, E).copy())}
    return out


def fit_many(x, task_L, W, epochs=1000, canonical=False, **kw):
    """W f32 [n][2*E*Lmax] packed (We then Wd with each task's own L), updated in place; all cores."""
    x = np.ascontiguousarray(x, np.float32)
    T, E = x.shape
    n = len(task_L)
    cfg = _cfg(E, 0, T, epochs, **kw)
    tl = np.ascontiguousarray(task_L, np.int32)
    loss, val = np.zeros((n, epochs), np.float32), np.zeros((n, epochs), np.float32)
    ep = np.zeros(n, np.int32)
    v = lambda a: a.ctypes.data_as(C.c_void_p)
    (_lib().fit_many_canonical if canonical else _lib().fit_many_fast)(C.byref(cfg), n, v(tl), v(x), v(W), W.shape[1], v(loss), v(val), v(ep))
    return loss, val, ep
