"""GPU test of hfrep_fit_windows_static, the rolling-refit entry point for hosts that cannot allocate device
memory (their own arena already holds the GPU, or an allocation's device-wide stall is not acceptable).  It is
called straight through the C ABI with caller-owned inputs and outputs and must reproduce hfrep_fit_windows —
which takes the scratch as a workspace argument — on two consecutive calls with different inputs, so that a
stale block left in the built-in array by the first call would show up in the second.  Kept in a file of its
own that pytest collects last."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(hf, static, R, starts, T_win, E, task_L, task_win, W0, epochs):
    lib, ptr = hf._lib.lib, hf._lib.ptr
    dev = R.device
    n, n_win = len(task_L), len(starts)
    n_fit = int(np.floor(T_win * .75))
    cfg = hf._lib.fit_cfg(E=E, T=T_win, n_fit=n_fit, L_max=int(max(task_L)), max_epochs=epochs, early_stop=0,
                          warps_per_task=3)
    sched = hf.engine.nadam_schedule(cfg, epochs * ((n_fit + cfg.batch - 1) // cfg.batch), dev)
    tL = torch.tensor(task_L, dtype=torch.int32, device=dev)
    tw = torch.tensor(task_win, dtype=torch.int32, device=dev)
    st = torch.tensor(starts, dtype=torch.int32, device=dev)
    out = dict(W=torch.zeros_like(W0), loss=torch.zeros((n, epochs), device=dev), val=torch.zeros((n, epochs), device=dev),
               epochs=torch.zeros(n, dtype=torch.int32, device=dev), status=torch.zeros(n, dtype=torch.int32, device=dev))
    stream = hf._lib.stream_ptr()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]            # every torch allocation of this call is done by now
    if static:
        rc = lib.hfrep_fit_windows_static(C.byref(cfg), ptr(R), R.shape[0], R.shape[1], 0, ptr(st), n_win, n, ptr(tL), ptr(tw),
                                          ptr(W0), W0.shape[1], None, 0, None, ptr(sched), ptr(out["W"]), ptr(out["loss"]),
                                          ptr(out["val"]), ptr(out["epochs"]), ptr(out["status"]), stream)
    else:
        ws_bytes = lib.hfrep_fit_windows_workspace(n_win, T_win, E)
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=dev)
        rc = lib.hfrep_fit_windows(C.byref(cfg), ptr(R), R.shape[0], R.shape[1], 0, ptr(st), n_win, n, ptr(tL), ptr(tw),
                                   ptr(W0), W0.shape[1], None, 0, None, ptr(sched), ptr(ws), ws_bytes, ptr(out["W"]),
                                   ptr(out["loss"]), ptr(out["val"]), ptr(out["epochs"]), ptr(out["status"]), stream)
    torch.cuda.synchronize()
    out["free_delta"] = free_before - torch.cuda.mem_get_info(dev)[0] if static else None
    return rc, out


def test_fit_windows_static_matches_workspace_variant():
    import hfrep as hf
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    rng = np.random.default_rng(11)
    E, L, T_win, epochs = 22, 5, 48, 20
    for call, T_all in enumerate((90, 75)):                 # two calls in a row, different data and window count
        R = torch.from_numpy((rng.normal(size=(T_all, E)) * .03).astype(np.float32)).to(dev)
        starts = list(range(0, T_all - T_win, 3))
        task_win = [w for w in range(len(starts)) for _ in range(2)]
        task_L = [L] * len(task_win)
        W0 = torch.from_numpy(hf.engine.pack_weights(
            [hf.engine.initial_weights(100 * call + i, E, L) for i in range(len(task_win))], E, 2 * E * L)).to(dev)
        rc_ref, ref = _call(hf, False, R, starts, T_win, E, task_L, task_win, W0, epochs)
        rc_new, new = _call(hf, True, R, starts, T_win, E, task_L, task_win, W0, epochs)
        assert rc_ref == 0 and rc_new == 0, hf._lib.lib.hfrep_last_error()
        assert int(new["status"].abs().sum()) == 0 and int(new["epochs"].min()) == epochs
        if call == 1:      # once the first call has loaded the array and the kernels, a call takes no device memory at all
            assert new["free_delta"] == 0, new["free_delta"]
        for k in ("W", "loss", "val"):                      # same kernels on the same blocks: 1e-5 relative is the suite's bar
            a, b = new[k].cpu().numpy(), ref[k].cpu().numpy()
            assert np.max(np.abs(a - b)) <= 1e-5 * np.max(np.abs(b)), (call, k)


def test_fit_windows_static_rejects_oversize():
    import hfrep as hf
    dev = torch.device("cuda:0")
    E, T_win = 22, 168
    n_win = hf._lib.STATIC_SCRATCH_BYTES // (4 * T_win * E) + 1       # one window too many
    cfg = hf._lib.fit_cfg(E=E, T=T_win, n_fit=126, L_max=4)
    one = torch.zeros(16, device=dev)
    p = hf._lib.ptr(one)
    rc = hf._lib.lib.hfrep_fit_windows_static(C.byref(cfg), p, 400, E, 0, p, n_win, 1, p, p, p, 2 * E * 4, None, 0, None, p,
                                              p, p, p, p, p, None)
    assert rc == -1
    msg = hf._lib.lib.hfrep_last_error().decode()
    assert "scratch" in msg and str(hf._lib.STATIC_SCRATCH_BYTES) in msg
