"""CPU tests of the host logic and of the C-ABI library's surface (no GPU compute calls)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, load_golden

import hfrep
from hfrep import _lib, engine
from oracle import ae_oracle, fitcore


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "hfrep.h")).read()
    declared = sorted(set(re.findall(r"\b(hfrep_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 13
    lib = ctypes.CDLL(hfrep.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/hfrep.h but not exported"
    assert sorted(_lib.EXPORTS) == declared
    assert _lib.lib.hfrep_version() >= 100


def test_struct_layouts_match_header_defaults():
    f = _lib.fit_cfg()
    assert (f.E, f.L_max, f.T, f.n_fit, f.batch, f.max_epochs, f.patience, f.early_stop) == (22, 8, 168, 126, 48, 1000, 5, 1)
    assert np.allclose([f.alpha, f.lr, f.beta_1, f.beta_2, f.epsilon, f.schedule_decay, f.decay_base],
                       [.2, 1e-3, .9, .999, 1e-7, .004, .96])
    assert (f.n_x, f.warps_per_task, f.dump_step) == (1, 0, 0)
    e = _lib.eval_cfg()
    assert (e.E, e.F, e.T_train, e.T_test, e.window, e.beta_rolling, e.want_oos) == (22, 13, 168, 169, 24, 0, 1)
    assert abs(e.alpha - .2) < 1e-7 and e.cost_param == .05 and e.phi == .5


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libhfrep.so"))
    with pytest.raises(_lib.HfrepError, match="no CPU fallback"):
        _lib._load()


def test_no_gpu_means_error_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.HfrepError, match="no CPU fallback"):
        hfrep.ReplicationData(np.zeros((40, 22)), np.zeros((40, 13)), np.zeros(40))


def test_window_tables_bit_exact():
    """SURVEY App. C, against the python ranges of the reference loops."""
    s, l = engine.window_index("rolling", 169, 24)           # Autoencoder_encapsulate.py:148
    assert s.dtype == np.int32 and np.array_equal(s, np.arange(145)) and np.all(l == 24)
    s, l = engine.window_index("expanding", 169, 24)         # :117
    assert np.array_equal(l, np.arange(2, 169)) and np.all(s == 0) and len(l) == 167
    s, l = engine.window_index("expost", 168, 24)            # helper.py:120
    assert np.array_equal(s, np.arange(1, 144)) and len(s) == 143
    assert len(engine.window_index("rolling", 10, 24)[0]) == 0     # window longer than the block
    assert engine.split_rows(337) == (168, 169)              # autoencoder_v4.ipynb:112
    with pytest.raises(_lib.HfrepError):
        _lib.check(_lib.lib.hfrep_window_index(9, 10, 2, None, None, 0))


def test_host_prep_matches_oracle(ref_inputs):
    x = ref_inputs["etf"][:168]
    assert np.array_equal(engine.minmax_scale(x), ae_oracle.minmax_fit_transform(x))
    assert np.array_equal(engine.minmax_scale(x), load_golden("ref_modeR_L2")["x_train_scaled"])   # sklearn's own output
    const = np.ones((5, 3)); const[:, 1] = [1, 2, 3, 4, 5]
    assert np.array_equal(engine.minmax_scale(const)[:, 0], np.zeros(5))      # zero range -> scale 1
    for L in (1, 8, 21):
        a, b = engine.initial_weights(123, 22, L), fitcore.initial_weights(123, 22, L)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert np.abs(a[0]).max() <= np.sqrt(6 / (22 + L))
    assert np.array_equal(engine.epoch_permutations(7, 5, 126), fitcore.make_perm(7, 5, 126))
    # device-side shuffle streams: host replica == oracle restatement, and every row is a permutation
    for seed in (0, 5, 123456789, 0xFFFFFFFF):
        tab = engine.device_permutations(seed, 4, 126)
        assert np.array_equal(tab, fitcore.device_perm(seed, 4, 126))
        assert all(sorted(r.tolist()) == list(range(126)) for r in tab)
    assert not np.array_equal(engine.device_permutations(5, 2, 126)[0], engine.device_permutations(5, 2, 126)[1])
    We, Wd = engine.initial_weights(1, 22, 5)
    row = engine.pack_weights([(We, Wd)], 22, 2 * 22 * 8)[0]
    We2, Wd2 = engine.unpack_weights(row, 22, 5)
    assert np.array_equal(We, We2) and np.array_equal(Wd, Wd2) and np.all(row[2 * 22 * 5:] == 0)


def test_nadam_schedule_matches_oracle():
    cfg = _lib.fit_cfg()
    sched = np.zeros((3000, _lib.SCHED_STRIDE), np.float32)
    _lib.check(_lib.lib.hfrep_nadam_schedule(ctypes.byref(cfg), 3000, _lib.ptr(sched)))
    opt = fitcore.Nadam(fitcore.FitConfig(), [(1,)])
    one = np.float32(1.)
    for t in range(3000):
        mu_t, mu_t1, ms_new, ms_next, vden = opt.coefficients()
        want = np.array([one - ms_new, one - ms_next, vden, one - mu_t, mu_t1, mu_t, 0, 0], np.float32)
        assert np.array_equal(sched[t], want), t          # the very scalars the oracle divides and multiplies by: bit-equal
        opt.m_cache = ms_new
        opt.t += 1


def test_argument_validation_without_gpu():
    """Bad shapes are rejected before any CUDA call."""
    cfg = _lib.fit_cfg(E=30)
    one = ctypes.c_void_p(16)
    rc = _lib.lib.hfrep_fit_batched(ctypes.byref(cfg), one, 1, one, None, one, 2000, None, 0, None, one, one, one, one, one, one,
                                    None, None)
    assert rc == -1 and b"E=30" in _lib.lib.hfrep_last_error()
    ecfg = _lib.eval_cfg(window=100)
    rc = _lib.lib.hfrep_predict_metrics(ctypes.byref(ecfg), one, one, one, one, one, 1, one, None, one, 400, *([None] * 11))
    assert rc == -1 and b"window" in _lib.lib.hfrep_last_error()


def test_shard_partition():
    from hfrep import distributed as D
    for n, w in [(21, 2), (1344, 8), (5, 8), (0, 4)]:
        parts = [D.shard(n, r, w) for r in range(w)]
        allidx = np.sort(np.concatenate(parts)) if n else np.zeros(0)
        assert np.array_equal(allidx, np.arange(n))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_gather_world_size_2_gloo(tmp_path):
    """The one exchange step of the multi-GPU path, on CPU with gloo and two processes."""
    script = tmp_path / "w.py"
    script.write_text(f"""
import os, sys
sys.path.insert(0, {ROOT!r})
import numpy as np, torch, torch.distributed as dist
from hfrep import distributed as D
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{{}}".format(os.environ["PORT"]),
                        rank=int(os.environ["RANK"]), world_size=2)
rank, world = D.world()
n = 21
mine = D.shard(n, rank, world)
local = torch.tensor([[float(i), float(i) * 2] for i in mine], dtype=torch.float64).reshape(len(mine), 2)
full = D.gather_results(local, n)
assert torch.equal(full[:, 0], torch.arange(n, dtype=torch.float64)), full
assert torch.equal(full[:, 1], 2 * torch.arange(n, dtype=torch.float64))
# runs of consecutive tasks (rolling windows): ragged too
mine_b = D.shard_blocks(n, rank, world)
assert len(mine_b) == (11 if rank == 0 else 10) and (np.diff(mine_b) == 1).all()
full_b = D.gather_results(torch.from_numpy(mine_b).to(torch.float64)[:, None], n, order="blocks")
assert torch.equal(full_b[:, 0], torch.arange(n, dtype=torch.float64)), full_b
# the packed-record gather: several tensors of different dtypes and shapes (odd byte counts, a missing one) in ONE collective
idx = torch.from_numpy(mine)
fields = [idx.to(torch.float64)[:, None, None] * torch.ones(1, 3, 2, dtype=torch.float64), None,
          idx.to(torch.int32), (idx % 7).to(torch.uint8)[:, None].repeat(1, 5), idx.to(torch.float32)[:, None] * torch.arange(3.)]
calls = []
orig = dist.all_gather_into_tensor
dist.all_gather_into_tensor = lambda *a, **k: (calls.append(1), orig(*a, **k))[1]
got = D.gather_records(fields, n)
dist.all_gather_into_tensor = orig
g = torch.arange(n)
assert len(calls) == 1 and got[1] is None
assert torch.equal(got[0], g.to(torch.float64)[:, None, None] * torch.ones(1, 3, 2, dtype=torch.float64))
assert torch.equal(got[2], g.to(torch.int32)) and got[2].dtype == torch.int32
assert torch.equal(got[3], (g % 7).to(torch.uint8)[:, None].repeat(1, 5))
assert torch.equal(got[4], g.to(torch.float32)[:, None] * torch.arange(3.))
dist.barrier(); dist.destroy_process_group()
print("ok", rank)
""")
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(os.environ, RANK=str(r), PORT=str(port)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for r in range(2)]
    outs = [p.communicate(timeout=180)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("ok" in o for o in outs)


def test_dropin_modules_resolve_to_the_package():
    code = ("import sys, hfrep; sys.path.insert(0, hfrep.DROPIN_PATH); import Autoencoder_encapsulate as A, helper as H; "
            "assert A.AE is hfrep.AE and H.random_sampling is hfrep.helper.random_sampling; "
            "import inspect; print(inspect.signature(A.AE.ante), inspect.signature(A.AE.train), inspect.signature(H.price_impact))")
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "(self, rf, hfd, window=24)" in out.stdout and "(self, patience=5, verbose=2, plot=True)" in out.stdout
    assert "(old_x, new_x, covMatrix, param=0.05, phi=0.5)" in out.stdout



def test_ff_alpha_grs_hk_match_the_notebooks_printed_table(ref_inputs):
    """FF3/FF5 alphas, GRS and HK tests (restated from the notebook's statsmodels call and R functions) against the
    values the reference printed for the 13 real indices (autoencoder_v4.ipynb, cell `hfd_res`): host code, no GPU."""
    import hfrep
    ff = load_golden("ref_ff_factors")
    hfd, etf = ref_inputs["hfd"][-144:], ref_inputs["etf"][-144:]
    h = hfrep.helper
    got = {"FF3F_alpha": [h.OLS_alpha(hfd[:, j], ff["three_factor"]) for j in range(13)],
           "FF5F_alpha": [h.OLS_alpha(hfd[:, j], ff["five_factor"]) for j in range(13)],
           "GRS_testF": [h.grstest(hfd[:, j], etf)[0][0] for j in range(13)],
           "GRS_test_pval": [h.grstest(hfd[:, j], etf)[1][0] for j in range(13)],
           "HK_testF": [h.hktest(hfd[:, j], etf)[0][0] for j in range(13)],
           "HK_test_pval": [h.hktest(hfd[:, j], etf)[1][0] for j in range(13)]}
    for k, v in got.items():
        assert np.max(np.abs(np.asarray(v, float).reshape(-1) - ff["want_" + k])) < 1e-6, k
    # several test assets at once (the R functions' general case): finite, p in [0, 1]
    hk, grs = h.hktest(hfd[:, :3], etf), h.grstest(hfd[:, :3], etf)
    assert hk.shape == (2, 1) and 0 <= hk[1][0] <= 1 and np.isfinite(grs).all()
