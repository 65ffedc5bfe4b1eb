"""CPU tests of the oracle itself (test infrastructure): the portable restatement against the
golden fixtures produced by the reference's own files, the deterministic known-answer table of
the notebook, and the Nadam restatement against torch.optim.NAdam."""
import os

import numpy as np
import pytest

from conftest import load_golden, rel_err
from oracle import ae_oracle, fitcore, refbridge


@pytest.mark.parametrize("name", ["ref_modeR_L2", "ref_modeR_L8_shuffled"])
def test_portable_oracle_matches_reference_run(ref_inputs, name):
    g = load_golden(name)
    etf, hfd, rf = ref_inputs["etf"], ref_inputs["hfd"], ref_inputs["rf"]
    L = g["We0"].shape[1]
    perm = fitcore.make_perm(int(g["perm_seed"]), 1000, 126) if "perm_seed" in g else None
    o = ae_oracle.run_task(etf[:168], hfd[:168], etf[168:], hfd[168:], rf, etf, L, g["We0"], g["Wd0"], perm=perm)
    assert o["epochs_run"] == int(g["epochs_run"])
    for k in ("We", "Wd", "loss", "val_loss"):
        assert np.array_equal(o[k], g[k]), k                 # same float32 arithmetic -> bit-identical
    for k in ("is_r2", "is_rmse", "oos_r2", "oos_rmse", "ante", "post", "turnover"):
        assert rel_err(o[k], g[k]) < 1e-10, k


@pytest.mark.parametrize("name", ["ref_modeR_L8_deadunit", "ref_modeR_L8_rank6"])
def test_portable_oracle_rank_deficient_windows(ref_inputs, name):
    """Rank-deficient OLS windows (dead / duplicated latent unit): the portable restatement against the reference's own
    ante / post / turnover (minimum-norm pinv solution, Autoencoder_encapsulate.py:151)."""
    g = load_golden(name)
    etf, hfd, rf = ref_inputs["etf"], ref_inputs["hfd"], ref_inputs["rf"]
    ae = ae_oracle.AEOracle(etf[:168], hfd[:168], etf[168:], hfd[168:], g["We"].shape[1])
    ae.We, ae.Wd = g["We"], g["Wd"]
    assert rel_err(ae.ante(rf), g["ante"]) < 1e-9
    assert rel_err(ae.weights, g["weights"]) < 1e-9
    assert rel_err(ae.post(etf), g["post"]) < 1e-9 and rel_err(ae.turnover(), g["turnover"]) < 1e-9
    assert np.linalg.matrix_rank(ae.latent[:24].astype(np.float64)) < g["We"].shape[1]


def test_portable_oracle_stacked_rows_run(ref_inputs):
    """The GAN-augmented fit of autoencoder_v4.ipynb:1368-1427 (1 848 stacked training rows): the portable restatement
    against the reference's own run on GAN/generated_data2022-07-09.pkl."""
    from sklearn.preprocessing import MinMaxScaler
    g = load_golden("ref_stacked_L8")
    etf, hfd, rf = ref_inputs["etf"], ref_inputs["hfd"], ref_inputs["rf"]
    sc = MinMaxScaler().fit(np.hstack([etf, hfd, rf.reshape(len(etf), -1)]))
    ret_gen = np.array([sc.inverse_transform(p) for p in g["gen_data"]])
    fa, rest = ae_oracle.factor_hf_split(ret_gen, 22)
    ae = ae_oracle.AEOracle(np.vstack([fa, etf[:168]]), np.vstack([rest[:, :13], hfd[:168]]), etf[168:], hfd[168:], 8)
    o = ae.train(g["We0"], g["Wd0"])
    assert o["epochs_run"] == int(g["epochs_run"]) and fitcore.split_at(1848, .25) == 1386
    for k in ("We", "Wd", "loss", "val_loss"):
        assert np.array_equal(o[k], g[k]), k
    assert abs(ae.model_IS_r2() - g["is_r2"]) < 1e-10
    assert rel_err(ae.ante(rf), g["ante"]) < 1e-10 and rel_err(ae.post(etf), g["post"]) < 1e-10


def test_generator_restatement_against_the_references_stored_paths():
    """oracle/gen_oracle.py (Keras LSTM / LeakyReLU / LayerNormalization / Dense restated) with the reference's saved
    generator weights on fresh N(0, 1) noise against GAN/generated_data2022-07-09.pkl — what the reference's own Keras run
    of that generator produced (autoencoder_v4.ipynb:1291; its noise was not saved, so only distributions can agree):
    overall mean / spread, every column's mean, the range."""
    import hfrep
    from oracle import gen_oracle
    layers = hfrep.generator.layers_from_fixture(np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_generator.npz")))
    ref = load_golden("ref_stacked_L8")["gen_data"]                       # (10, 168, 36) float32
    out = gen_oracle.forward(layers, np.random.default_rng(0).normal(0, 1, (40, 168, 36)))
    assert out.shape == (40, 168, 36) and out.dtype == np.float32
    assert abs(out.mean() - ref.mean()) < 5e-3 and abs(out.std() / ref.std() - 1) < .05
    assert np.max(np.abs(out.mean((0, 1)) - ref.mean((0, 1)))) < .015
    assert ref.min() - .15 < out.min() < 0 and 1 < out.max() < ref.max() + .15


@pytest.mark.skipif(not refbridge.available(), reason="reference checkout not present on this machine")
def test_hdf5_reader_on_the_references_model_files():
    """hfrep.h5lite (no h5py here) on every generator the reference saved: groups, attributes (fixed and variable-length
    strings), float32 datasets; the fixture the GPU tests use equals the file's content."""
    import glob
    import json
    import hfrep
    files = sorted(glob.glob(os.path.join(refbridge.REFERENCE_ROOT, "GAN", "trained_generator", "**", "*.h5"), recursive=True))
    assert len(files) == 8
    for path in files:
        f = hfrep.h5lite.File(path)
        a = f.root.attrs
        assert a["keras_version"] == "2.7.0" and a["backend"] == "tensorflow"
        cfg = json.loads(a["model_config"])
        assert cfg["class_name"] in ("Functional", "Sequential")
        data = f.root["model_weights"].datasets()
        assert data and all(v.dtype == np.float32 and np.isfinite(v).all() for v in data.values())
    layers = hfrep.generator.read_keras_model(os.path.join(refbridge.REFERENCE_ROOT, "GAN", "trained_generator",
                                                            "MTTS_GAN_GP20220621_02-49-32.h5"))
    fix = hfrep.generator.layers_from_fixture(np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_generator.npz")))
    assert [c for c, _, _ in layers] == [c for c, _, _ in fix]
    for (_, _, a), (_, _, b) in zip(layers, fix):
        assert len(a) == len(b) and all(np.array_equal(x, y) for x, y in zip(a, b))
    assert [w.shape for w in layers[1][2]] == [(36, 400), (100, 400), (400,)]       # LSTM(100) on 36 features


def test_structural_constants(ref_inputs):
    """Window bookkeeping pinned by the notebook."""
    g = load_golden("ref_modeR_L2")
    assert ref_inputs["etf"].shape == (337, 22) and ref_inputs["hfd"].shape == (337, 13)
    assert g["x_train_scaled"].shape == (168, 22)            # 168/169 split, autoencoder_v4.ipynb:112
    assert g["oos_r2"].shape == (167,)                       # 167 expanding prefixes, :319
    assert g["ante"].shape == (144, 13) and str(g["ante_index0"]) == "2010-05-31"   # :929
    assert fitcore.split_at(168, .25) == 126 and fitcore.split_at(1848, .25) == 1386


def test_known_answer_table(ref_inputs):
    """hfd_res (autoencoder_v4.ipynb:1015): deterministic statistics of the real indices, last 144 months."""
    hfd, rf = ref_inputs["hfd"][-144:], ref_inputs["rf"][-144:]
    sharpe = [ae_oracle.annualized_sharpe_ratio(hfd[:, f], rf) for f in range(13)]
    expect = [0.725028, 0.763790, 0.390113, 0.164249, 0.372265, 0.578300, 0.287477, 0.593060, 1.183535,
              0.932520, 0.541682, 0.214612, 1.204837]
    assert np.allclose(sharpe, expect, atol=5e-7)
    h = hfd[:, 0]
    got = [ae_oracle.omega_ratio(h, 0), ae_oracle.omega_ratio(h, .1), ae_oracle.historical_cvar(h),
           ae_oracle.ceq(h, rf, 2), ae_oracle.skewness(h), ae_oracle.kurtosis(h)]
    assert np.allclose(got, [1.908581, 0.593268, -0.031864, 0.033211, -1.321605, 6.283068], atol=5e-7)


def test_helper_restatements_match_reference():
    g = load_golden("ref_helpers")
    assert np.allclose(ae_oracle.normalization(g["norm_Y"], g["norm_X"], g["norm_beta"], 24), g["norm_out"], rtol=1e-13)
    cov = np.cov(g["tc_window"], rowvar=False, ddof=1)
    assert np.allclose(ae_oracle.transaction_cost(g["tc_old"], g["tc_new"], cov), g["tc_out"], rtol=1e-12)
    assert np.allclose(ae_oracle.price_impact(g["tc_old"], g["tc_new"], cov), g["pi_out"], rtol=1e-12)
    fa, hf = ae_oracle.factor_hf_split(g["split_in"], 22)
    assert np.array_equal(fa, g["split_factor"]) and np.array_equal(hf, g["split_hf"])
    assert np.array_equal(ae_oracle.random_window_starts(337, 16, 48, 123), g["rs_starts"])
    assert np.array_equal(ae_oracle.random_sampling(g["rs_in"], g["rs_starts"], 48), g["rs_out"])
    post = ae_oracle.ex_post_return(g["epr_ante"], int(g["epr_window"]), g["epr_strat"], g["epr_etf"])
    assert rel_err(post, g["epr_out"]) < 1e-12
    assert np.array_equal(np.transpose(g["cab_in"], (2, 0, 1)), g["cab_out"])     # reshape_cab is a transpose


def test_nadam_against_torch():
    """torch.optim.NAdam(eps=1e-7, momentum_decay=4e-3) is formula-identical to Keras 2.7 Nadam (App. A.4)."""
    import torch
    rng = np.random.default_rng(0)
    w0 = rng.normal(size=(22, 8)).astype(np.float32)
    grads = [rng.normal(size=(22, 8)).astype(np.float32) * .01 for _ in range(200)]
    w = w0.copy()
    opt = fitcore.Nadam(fitcore.FitConfig(), [w.shape])
    for gr in grads:
        opt.step([w], [gr])
    tw = torch.tensor(w0.astype(np.float64), requires_grad=True)
    topt = torch.optim.NAdam([tw], lr=1e-3, betas=(.9, .999), eps=1e-7, momentum_decay=4e-3)
    for gr in grads:
        tw.grad = torch.tensor(gr.astype(np.float64))
        topt.step()
    assert rel_err(w, tw.detach().numpy()) < 2e-6


def test_early_stopping_rule():
    """App. A.6: strict '<', wait counted before the comparison, never stops at epoch 0."""
    rng = np.random.default_rng(1)
    x = rng.uniform(0, 1, (64, 6))
    We0, Wd0 = fitcore.initial_weights(3, 6, 2)
    perm = fitcore.make_perm(9, 400, 48)
    out = fitcore.fit(x, We0, Wd0, fitcore.FitConfig(epochs=400, batch_size=16, patience=2), perm=perm)
    v = out["val_loss"]
    n = out["epochs_run"]
    assert n == len(v) == len(out["loss"])
    if n < 400:
        best_before = np.min(v[:n - 2])
        assert not (v[n - 2] < best_before) and not (v[n - 1] < min(best_before, v[n - 2]))
    out0 = fitcore.fit(x, We0, Wd0, fitcore.FitConfig(epochs=5, batch_size=16, patience=0))
    assert out0["epochs_run"] == 2          # patience 0 still needs epoch > 0


@pytest.mark.skipif(not refbridge.available(), reason="reference checkout not present on this machine")
def test_reference_import_is_unmodified_source():
    ae_mod, helper, keras = refbridge.load()
    try:
        assert ae_mod.__file__ == os.path.join(refbridge.REFERENCE_ROOT, "Autoencoder_encapsulate.py")
        assert helper.__file__ == os.path.join(refbridge.REFERENCE_ROOT, "helper.py")
        g = load_golden("ref_helpers")
        assert np.allclose(helper.normalization(g["norm_Y"], g["norm_X"], g["norm_beta"], 24), g["norm_out"])
    finally:
        refbridge.unload()


def test_fused_multiply_add_emulation():
    """fitcore.fma32 (float64 emulation with the double-rounding repair) against libm's fmaf, on cancelling operands."""
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    libm.fmaf.restype = ctypes.c_float
    libm.fmaf.argtypes = [ctypes.c_float] * 3
    rng = np.random.default_rng(0)
    n = 20000
    a, b = rng.normal(size=n).astype(np.float32), rng.normal(size=n).astype(np.float32)
    c = (-(a.astype(np.float64) * b) * (1 + rng.normal(size=n) * 1e-4)).astype(np.float32)
    c[::7] = rng.normal(size=len(c[::7])).astype(np.float32) * 1e3
    want = np.array([libm.fmaf(float(x), float(y), float(z)) for x, y, z in zip(a, b, c)], np.float32)
    assert np.array_equal(fitcore.fma32(a, b, c), want)


@pytest.mark.parametrize("E,L,T,batch,epochs", [(22, 8, 168, 48, 10), (7, 13, 90, 25, 8), (10, 3, 64, 16, 8), (24, 24, 100, 32, 4),
                                                (22, 1, 57, 5, 4)])
def test_canonical_order_numpy_equals_c(E, L, T, batch, epochs):
    """The canonical summation order written twice — numpy (emulated fused multiply-adds) and C (fmaf) — gives the same
    bits: kernels, both histories, gradients.  Ragged shapes: partial batches, batches shorter than the six row classes."""
    rng = np.random.default_rng(E * 31 + L)
    x = rng.uniform(0, 1, (T, E))
    We0, Wd0 = fitcore.initial_weights(5, E, L)
    perm = fitcore.make_perm(3, epochs, fitcore.split_at(T, .25))
    cfg = fitcore.FitConfig(epochs=epochs, batch_size=batch, early_stop=False)
    a = fitcore.fit(x, We0, Wd0, cfg, perm=perm, backend="numpy", record_grads_at=(2,))
    b = fitcore.fit(x, We0, Wd0, cfg, perm=perm, backend="c", record_grads_at=(2,))
    for k in ("We", "Wd", "loss", "val_loss"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["grads"][2][0], b["grads"][2][0]) and np.array_equal(a["grads"][2][1], b["grads"][2][1])


def test_c_restatement_on_the_reference_runs():
    """The C restatement in canonical order reproduces the fixtures bit for bit (they were written by the numpy path:
    make_golden records gradients at three steps, which fitcore only does in numpy), deterministic order for 1000 epochs
    and shuffled order with the same early stop; in its natural loop order it agrees within the flip bound below."""
    from oracle import c_oracle
    g = load_golden("ref_modeR_L21")
    o = c_oracle.fit(g["x_train_scaled"], g["We0"], g["Wd0"], epochs=1000)
    assert o["epochs_run"] == 1000
    for k in ("We", "Wd", "loss", "val_loss"):
        assert np.array_equal(o[k], g[k]), k
    gs = load_golden("ref_modeR_L8_shuffled")
    perm = fitcore.make_perm(int(gs["perm_seed"]), 1000, 126)
    o = c_oracle.fit(gs["x_train_scaled"], gs["We0"], gs["Wd0"], epochs=1000, perm=perm)
    assert o["epochs_run"] == int(gs["epochs_run"]) and np.array_equal(o["We"], gs["We"]) and np.array_equal(o["val_loss"], gs["val_loss"])
    o = c_oracle.fit(gs["x_train_scaled"], gs["We0"], gs["Wd0"], epochs=1000, perm=perm, canonical=False)
    assert o["epochs_run"] == int(gs["epochs_run"])
    assert np.max(np.abs(o["val_loss"] - gs["val_loss"]) / gs["val_loss"]) < 1e-5


@pytest.mark.parametrize("name", ["ref_modeR_L2", "ref_modeR_L8", "ref_modeR_L21", "ref_modeR_L8_shuffled"])
def test_fit_restatement_against_independent_torch_fit(name):
    """oracle/fitcore.py (the arbiter of every fit parity test) against oracle/torch_fit.py — torch autograd, torch's
    leaky_relu / mse_loss and torch.optim.NAdam, nothing of it written here — over the full reference runs: the same
    split, batches, loss reduction, LeakyReLU gradient, Nadam and early-stop bookkeeping must come out.
    Bound: torch sums in another order, so histories agree to ~1e-6 — except after a LeakyReLU pre-activation has passed
    within rounding of zero (fitcore's module docstring: one row's gradient changes five-fold for one step), which shows
    as a few isolated epochs off by up to ~7e-5.  Checked: stop epoch equal; median deviation < 5e-6; at most 1 % of the
    epochs beyond 1e-5 and none beyond 2e-4; final kernels within 5e-5."""
    from oracle import torch_fit
    g = load_golden(name)
    perm = fitcore.make_perm(int(g["perm_seed"]), 1000, 126) if "perm_seed" in g else None
    o = torch_fit.fit(g["x_train_scaled"], g["We0"], g["Wd0"], fitcore.FitConfig(), perm=perm)
    n = int(g["epochs_run"])
    assert o["epochs_run"] == n
    for k in ("loss", "val_loss"):
        d = np.abs(o[k] - g[k]) / g[k]
        assert np.median(d) < 5e-6 and np.mean(d > 1e-5) <= .01 and d.max() < 2e-4, (k, np.median(d), d.max())
    assert rel_err(o["We"], g["We"]) < 5e-5 and rel_err(o["Wd"], g["Wd"]) < 5e-5


def test_no_validation_rows_never_stops_early():
    """validation_split = 0: there is no val_loss, Keras' EarlyStopping warns and never stops."""
    rng = np.random.default_rng(2)
    x = rng.uniform(0, 1, (40, 6))
    We0, Wd0 = fitcore.initial_weights(3, 6, 2)
    cfg = fitcore.FitConfig(epochs=12, batch_size=16, validation_split=0., patience=1)
    for backend in ("numpy", "c"):
        assert fitcore.fit(x, We0, Wd0, cfg, backend=backend)["epochs_run"] == 12
