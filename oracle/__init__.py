"""CPU oracle for the CP-PRE hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import it, and only as the checker (or as the timed
CPU baseline).  The product package ``cp_pre_b200`` never imports it and fails
loudly when its CUDA library is missing.

Contents
--------
``cpu_oracle``   NumPy restatement of the reference algorithms (stencil tables,
                 kernel embedding, zero-padded cross-correlation, boundary
                 padding, spectral path, residual assemblies, conformal helpers).
                 Every function cites the reference file:line it follows.
``torch_port``   The same path expressed with the library calls the reference
                 itself makes on CPU (``F.conv1d/2d/3d`` + elementwise); used as
                 the multi-threaded timed CPU baseline (kind = "port").
``synth``        Seeded synthetic field generators for the BASELINE configs.
``ref_shim``     Imports the unmodified reference from ``/root/reference``
                 (build container only -- that path does not exist on the GPU
                 box) so ``make_golden.py`` can pin the restatement.
``make_golden``  Writes ``tests/golden/*.npz`` from the live reference.

Parity status: the stencil / boundary / spectral / vector half is pinned against
outputs of the reference itself (``tests/golden``).  The five conformal helpers
live in an un-vendored submodule (Neural_PDE @ fdf1dcc3, ``UQ/inductive_cp.py``)
that is absent from the reference tree: ``modulation_func``, ``ncf_metric_joint``
and ``emp_cov_joint`` are pinned on their in-tree inline twins
(``Tests/test_advection_inv_sampling_marginal.py:428-431,465``,
``Joint/Advection_Residuals_CP.py:327-329``); ``calibrate`` and ``emp_cov`` have
no in-tree twin -- **parity unpinned** for those two (restated as standard split
conformal prediction, see ``cpu_oracle.calibrate``).
"""
