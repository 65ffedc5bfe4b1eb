"""CPU oracle for the replication hot path — TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import anything below this package.  The product package never
does: it fails loudly when its CUDA library is missing.

Parity status (also in DESIGN.md):

* The reference's own arithmetic (``Autoencoder_encapsulate.py:107-224``, ``helper.py:10-153``)
  is *pinned*: ``oracle/make_golden.py`` imports those two files unmodified from
  ``/root/reference`` over the shim packages in ``oracle/shims`` and writes their outputs to
  ``tests/golden``; ``oracle/ae_oracle.py`` (the portable restatement) is checked against those
  fixtures, and the deterministic ``hfd_res`` Sharpe table printed in
  ``autoencoder_v4.ipynb:1015`` is a known-answer test.
* The Keras / statsmodels layer underneath (Dense, LeakyReLU, MSE, Nadam, fit's split and
  batching, EarlyStopping, OLS) is **parity unpinned**: neither package can be installed here,
  the reference has no tests, so ``oracle/fitcore.py`` restates Keras 2.7 semantics from their
  published definition and is cross-checked only against
  ``torch.optim.NAdam``.
"""
