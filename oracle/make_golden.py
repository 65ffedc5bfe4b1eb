"""Generate tests/golden/*.npz by running the REFERENCE'S OWN FILES (unmodified, imported
from /root/reference over oracle/shims — see oracle/refbridge.py) on the reference's own
cleaned CSVs.  Runs only in the build container; the fixtures travel, the reference does not.

    python -m oracle.make_golden            # writes tests/golden/ref_modeR_L{2,8,21}.npz, ref_helpers.npz, ...

Inputs that the reference would draw from TensorFlow's RNG (initial kernels, shuffles) are
explicit and stored in the fixture, so the CUDA path and oracle/ae_oracle.py can be given the
very same ones.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import fitcore, refbridge  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def run_reference_AE(ae_mod, keras, data, L, seed, perm=None, patience=5, epochs=None, grads_at=(0, 1, 100)):
    etf, hfd, rf = data
    from sklearn.model_selection import train_test_split
    X_train, X_test, Y_train, Y_test = train_test_split(etf, hfd, shuffle=False, test_size=.5)   # ipynb:112
    We0, Wd0 = fitcore.initial_weights(seed, etf.shape[1], L)
    feed = iter([We0, Wd0])
    keras.CONTROL.update(w0=lambda i, o: next(feed), perm=perm, record_grads_at=tuple(grads_at),
                         fit_overrides=({} if epochs is None else {"epochs": epochs}))
    with refbridge.positional_series_getitem():
        ae = ae_mod.AE(X_train, Y_train, X_test, Y_test, L)                      # ipynb:159
        ae.train(patience=patience, verbose=0, plot=False)                       # ipynb:160
        fit = keras.CONTROL["last_fit"]
        res = dict(
            We0=We0, Wd0=Wd0, We=fit["We"], Wd=fit["Wd"], loss=fit["loss"], val_loss=fit["val_loss"],
            epochs_run=np.int32(fit["epochs_run"]), x_train_scaled=np.asarray(ae._x_train),
            is_r2=ae.model_IS_r2(), is_rmse=ae.model_IS_RMSE(),                  # ipynb:162-163
            oos_r2=np.array(ae.model_OOS_r2()), oos_rmse=np.array(ae.model_OOS_RMSE()),   # ipynb:164-165
        )
        for t, (gWe, gWd) in fit["grads"].items():
            res[f"grad_We_step{t}"], res[f"grad_Wd_step{t}"] = gWe, gWd
        ante = ae.ante(rf, hfd)                                                  # ipynb:899
        from cleaned_names import hfd_fullname
        turn = ae.turnover(hfd_fullname)                                         # ipynb:900
        post = ae.post(etf)                                                      # ipynb:901
        res.update(ante=ante.values, post=post.values, turnover=turn["Turnover"].values,
                   weights=np.array([w.values for w in ae.strat_weight_on_etf]),            # (144, 22, 13)
                   reshaped0=ae.reshape_strat_weight_on_etf[0].values,
                   latent=ae.autoencoder.encoder.predict(ae._x_test.values if hasattr(ae._x_test, "values") else ae._x_test),
                   ante_index0=str(ante.index[0].date()), ante_columns=np.array(list(ante.columns)))
    keras.CONTROL.update(w0=None, perm=None, record_grads_at=(), fit_overrides={})
    return res


def run_reference_deadunit(ae_mod, keras, data, g, dead=(3,), dup=None):
    """Rank-deficient OLS windows through the reference's own `ante`/`post`/`turnover` (Autoencoder_encapsulate.py:133-224):
    the trained L=8 kernels of fixture `g` with encoder column(s) `dead` zeroed (a latent unit that never fires) and,
    optionally, column dup[1] made a copy of column dup[0] (two identical latent factors).  `OLS(Y, X).fit()` (:151) then
    returns the MINIMUM-NORM solution (statsmodels' pinv)."""
    etf, hfd, rf = data
    from sklearn.model_selection import train_test_split
    X_train, X_test, Y_train, Y_test = train_test_split(etf, hfd, shuffle=False, test_size=.5)
    We, Wd = g["We"].copy(), g["Wd"].copy()
    for c in dead:
        We[:, c] = 0.
    if dup is not None:
        We[:, dup[1]] = We[:, dup[0]]
    feed = iter([We, Wd])
    keras.CONTROL.update(w0=lambda i, o: next(feed), perm=None, record_grads_at=(), fit_overrides={})
    with refbridge.positional_series_getitem():
        ae = ae_mod.AE(X_train, Y_train, X_test, Y_test, We.shape[1])
        ae.autoencoder = ae_mod.Autoencoder(We.shape[1])          # what train() builds at :78, with the given kernels
        ante = ae.ante(rf, hfd)
        from cleaned_names import hfd_fullname
        turn = ae.turnover(hfd_fullname)
        post = ae.post(etf)
        res = dict(We=We, Wd=Wd, ante=ante.values, post=post.values, turnover=turn["Turnover"].values,
                   weights=np.array([w.values for w in ae.strat_weight_on_etf]))
    keras.CONTROL.update(w0=None)
    return res


def run_reference_stacked(ae_mod, helper, keras, data, L, seed):
    """The notebook's GAN-augmented run (autoencoder_v4.ipynb cells 47-51, raw lines :1368-1427) on the reference's own
    generated paths GAN/generated_data2022-07-09.pkl: inverse MinMax with the scaler of the joined real data,
    helper.factor_hf_split, vstack on top of the real training half, then AE(...).train() — 1 848 training rows."""
    etf, hfd, rf = data
    from sklearn.model_selection import train_test_split
    from sklearn.preprocessing import MinMaxScaler
    gen_data = helper.dic_read(os.path.join(refbridge.REFERENCE_ROOT, "GAN", "generated_data2022-07-09.pkl"))
    data_scale = MinMaxScaler()
    data_scale.fit(etf.join(hfd).join(rf))                                              # cell 47
    ret_gen = np.array([data_scale.inverse_transform(gen_data[i]) for i in range(gen_data.shape[0])])
    train_factor, train_temp = helper.factor_hf_split(ret_gen, 22)                      # cell 48
    train_hf = train_temp[:, :13]
    X_train, X_test, Y_train, Y_test = train_test_split(etf, hfd, shuffle=False, test_size=.5)
    X_tr = np.vstack([train_factor, np.array(X_train)])                                 # cell 50
    Y_tr = np.vstack([train_hf, np.array(Y_train)])
    We0, Wd0 = fitcore.initial_weights(seed, etf.shape[1], L)
    feed = iter([We0, Wd0])
    keras.CONTROL.update(w0=lambda i, o: next(feed), perm=None, record_grads_at=(), fit_overrides={})
    with refbridge.positional_series_getitem():
        ae = ae_mod.AE(X_tr, Y_tr, X_test, Y_test, L)                                   # cell 51
        ae.train(verbose=0, plot=False)
        fit = keras.CONTROL["last_fit"]
        res = dict(gen_data=np.asarray(gen_data), We0=We0, Wd0=Wd0, We=fit["We"], Wd=fit["Wd"], loss=fit["loss"],
                   val_loss=fit["val_loss"], epochs_run=np.int32(fit["epochs_run"]), n_train_rows=np.int32(len(X_tr)),
                   is_r2=ae.model_IS_r2(), is_rmse=ae.model_IS_RMSE(),
                   oos_r2=np.array(ae.model_OOS_r2()), oos_rmse=np.array(ae.model_OOS_RMSE()))
        ante = ae.ante(rf, hfd)
        post = ae.post(etf)
        res.update(ante=ante.values, post=post.values)
    keras.CONTROL.update(w0=None)
    return res


def generator_fixture():
    """The reference's saved generator (GAN/trained_generator/MTTS_GAN_GP20220621_02-49-32.h5, the file the notebook loads at
    autoencoder_v4.ipynb:1277) read with the package's own HDF5 reader: layer list (the file's model_config JSON) and
    float32 weights, so the GPU box — which has neither the file nor h5py — can rebuild the model."""
    import json
    import importlib
    gen = importlib.import_module("hfrep").generator
    path = os.path.join(refbridge.REFERENCE_ROOT, "GAN", "trained_generator", "MTTS_GAN_GP20220621_02-49-32.h5")
    layers = gen.read_keras_model(path)
    out = {"layers_json": np.array(json.dumps([[c, cfg] for c, cfg, _ in layers]))}
    for i, (_, _, ws) in enumerate(layers):
        for j, w in enumerate(ws):
            out[f"w_{i}_{j}"] = w
    return out


def helper_vectors(helper):
    """Outputs of the reference's numeric helpers on seeded random inputs."""
    import random
    rng = np.random.default_rng(7)
    out = {}
    Y, X, beta = rng.normal(size=(24, 13)) * .02, rng.normal(size=(24, 5)) * .03, rng.normal(size=(5, 13))
    out.update(norm_Y=Y, norm_X=X, norm_beta=beta, norm_out=helper.normalization(Y, X, beta, 24))
    old, new = rng.normal(size=22), rng.normal(size=22)
    A = rng.normal(size=(24, 22)) * .04
    cov = pd.DataFrame(A).cov()
    out.update(tc_old=old, tc_new=new, tc_window=A, tc_out=helper.transaction_cost(old, new, cov),
               pi_out=helper.price_impact(old, new, cov))
    frames = [pd.DataFrame(rng.normal(size=(22, 13))) for _ in range(6)]
    out["cab_in"] = np.array([f.values for f in frames])
    out["cab_out"] = np.array([f.values for f in helper.reshape_cab(frames)])
    arr = rng.normal(size=(4, 168, 36)).astype(np.float32)
    fa, hf = helper.factor_hf_split(arr, 22)
    out.update(split_in=arr, split_factor=fa, split_hf=hf)
    data = rng.normal(size=(337, 35))
    random.seed(123)
    samp = helper.random_sampling(data, 16, 48)
    random.seed(123)
    starts = np.array([random.randint(0, 337 - 48) for _ in range(16)], np.int32)
    out.update(rs_in=data, rs_out=samp, rs_starts=starts)
    # ex_post_return on small synthetic inputs: T=10 months, window 6, 3 funds, 5 ETFs
    T, w, Fn, E = 10, 6, 3, 5
    ante = pd.DataFrame(rng.normal(size=(T, Fn)) * .01)
    strat = [pd.DataFrame(rng.normal(size=(T, E))) for _ in range(Fn)]
    etf = pd.DataFrame(rng.normal(size=(T + w, E)) * .03)
    with refbridge.positional_series_getitem():
        post = helper.ex_post_return(ante, w, strat, etf)
    out.update(epr_ante=ante.values, epr_strat=np.array([s.values for s in strat]), epr_etf=etf.values,
               epr_window=np.int32(w), epr_out=post.values)
    return out


# autoencoder_v4.ipynb, output of cell "hfd_res" (data_analysis(hfd[-144:], ..., span=factor_etf_data, real_data=True,
# start='2010-05-31', end='2022-04-30')): the columns the notebook computes through statsmodels and R, as printed there
# (6 decimals), rows in hfd column order.  The reference's own run; typed in because neither statsmodels nor R is here.
HFD_RES_REAL = {
    "FF3F_alpha": [0.000785, 0.001608, -0.000468, -0.000521, -0.000613, 0.001002, -0.001443, 0.001154, 0.002767, 0.003339,
                   -0.000749, 0.000360, 0.002785],
    "FF5F_alpha": [0.000820, 0.001615, -0.000447, -0.000518, -0.000564, 0.001045, -0.001386, 0.001171, 0.002788, 0.003381,
                   -0.000700, 0.000386, 0.002814],
    "GRS_testF": [7.392153, 8.236073, 2.162217, 1.759139, 1.452288, 9.067233, 0.130346, 7.380064, 25.902891, 8.431606,
                  2.458737, 0.121840, 20.653348],
    "HK_testF": [9.357224, 7.793611, 1.406071, 9.439554, 2.616191, 11.474257, 0.638452, 6.257770, 24.243047, 9.357745,
                 2.226949, 0.117562, 19.318581],
    "GRS_test_pval": [0.007514, 0.004848, 0.144036, 0.187230, 0.230513, 0.003169, 0.718703, 0.007562, 0.000001, 0.004384,
                      0.119484, 0.727654, 0.000013],
    "HK_test_pval": [0.000167, 0.000655, 0.249080, 0.000155, 0.077212, 0.000027, 0.529879, 0.002593, 0.000000, 0.000166,
                     0.112260, 0.889187, 0.000000],
}


def ff_factor_fixture():
    """The notebook's three_factor / five_factor cells run on the reference's own data files (pandas only), cut to the
    144 out-of-sample months, plus the printed table above."""
    out = {}
    for key, fname in (("three_factor", "F-F_Research_Data_Factors_daily.CSV"),
                       ("five_factor", "F-F_Research_Data_5_Factors_2x3_daily.CSV")):
        f = pd.read_csv(os.path.join(refbridge.REFERENCE_ROOT, "data", fname), usecols=['Date', 'Mkt-RF', 'SMB', 'HML'])
        f['Date'] = pd.to_datetime(f['Date'], format='%Y%m%d')
        f.set_index('Date', inplace=True)
        f = f.resample('ME').sum()
        f = np.log(f / 100 + 1)
        f = f.loc['1994-04-30':'2022-04-30', :]
        cut = f.loc['2010-05-31':'2022-04-30', ]
        out[key] = cut.values
        out[key + "_dates"] = np.array([str(d.date()) for d in cut.index])
    out.update({"want_" + k: np.array(v) for k, v in HFD_RES_REAL.items()})
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    ae_mod, helper, keras = refbridge.load()
    etf, hfd, rf = refbridge.read_reference_data()
    names = helper.dic_read(os.path.join(refbridge.REFERENCE_ROOT, "cleaned_data", "hfd_fullname.pkl"))
    mod = type(sys)("cleaned_names")
    mod.hfd_fullname = names
    sys.modules["cleaned_names"] = mod

    np.savez_compressed(os.path.join(OUT, "ref_inputs.npz"), etf=etf.values, hfd=hfd.values, rf=rf.values,
                        etf_columns=np.array(list(etf.columns)), hfd_columns=np.array(list(hfd.columns)),
                        dates=np.array([str(d.date()) for d in etf.index]),
                        hfd_fullname=np.array(list(names.values())))
    for L in (2, 8, 21):
        res = run_reference_AE(ae_mod, keras, (etf, hfd, rf), L, seed=123)
        np.savez_compressed(os.path.join(OUT, f"ref_modeR_L{L}.npz"), **res)
        print(f"L={L}: epochs_run={res['epochs_run']} is_r2={res['is_r2']:.4f} loss[-1]={res['loss'][-1]:.6g} "
              f"turnover[0]={res['turnover'][0]:.3f}")
    # shuffled order + early stopping: per-epoch permutations from the documented stream
    perm = fitcore.make_perm(123, 1000, 126)
    res = run_reference_AE(ae_mod, keras, (etf, hfd, rf), 8, seed=124, perm=perm, patience=5)
    res["perm_seed"] = np.int32(123)
    np.savez_compressed(os.path.join(OUT, "ref_modeR_L8_shuffled.npz"), **res)
    print(f"L=8 shuffled: epochs_run={res['epochs_run']} is_r2={res['is_r2']:.4f}")
    g8 = dict(np.load(os.path.join(OUT, "ref_modeR_L8.npz")))
    np.savez_compressed(os.path.join(OUT, "ref_modeR_L8_deadunit.npz"),
                        **run_reference_deadunit(ae_mod, keras, (etf, hfd, rf), g8, dead=(3,)))
    np.savez_compressed(os.path.join(OUT, "ref_modeR_L8_rank6.npz"),
                        **run_reference_deadunit(ae_mod, keras, (etf, hfd, rf), g8, dead=(3,), dup=(1, 6)))
    res = run_reference_stacked(ae_mod, helper, keras, (etf, hfd, rf), 8, seed=123)
    np.savez_compressed(os.path.join(OUT, "ref_stacked_L8.npz"), **res)
    print(f"stacked rows L=8: {res['n_train_rows']} rows, epochs_run={res['epochs_run']} is_r2={res['is_r2']:.4f}")
    np.savez_compressed(os.path.join(OUT, "ref_generator.npz"), **generator_fixture())
    np.savez_compressed(os.path.join(OUT, "ref_helpers.npz"), **helper_vectors(helper))
    np.savez_compressed(os.path.join(OUT, "ref_ff_factors.npz"), **ff_factor_fixture())
    refbridge.unload()


if __name__ == "__main__":
    main()
