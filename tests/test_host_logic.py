"""Host-side logic of the drop-in API and the C-ABI surface.  CPU only: nothing here launches
a kernel (the library is loaded, its symbols are checked, and its host-only entry points --
boundary index maps, TMA eligibility, error reporting -- are exercised)."""
import ast
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import cp_pre_b200
from conftest import ROOT, load_golden
from cp_pre_b200 import _lib as L
from cp_pre_b200 import ops
from oracle import cpu_oracle as O

cp_pre_b200.install()


def _mods():
    import ConvOps_0d
    import ConvOps_1d
    import ConvOps_Spatial
    import Utils.ConvOps_2d as ConvOps_2d
    return {"2d": ConvOps_2d, "1d": ConvOps_1d, "0d": ConvOps_0d, "spatial": ConvOps_Spatial}


def test_library_loads_and_exports_every_declared_symbol():
    lib = L.load()
    header = open(os.path.join(ROOT, "include", "cppre.h")).read()
    declared = set(re.findall(r"\b(cppre_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 19
    for name in declared:
        assert hasattr(lib, name), f"libcppre.so lacks {name}"
        assert name in L.SIGNATURES, f"_lib.py does not bind {name}"
    assert lib.cppre_version() == 100
    assert isinstance(L.launch_count(), int)


def test_abi_struct_layout_matches_header_constants():
    header = open(os.path.join(ROOT, "include", "cppre.h")).read()
    for name, val in (("CPPRE_MAX_FIELDS", L.MAX_FIELDS), ("CPPRE_MAX_SLOTS", L.MAX_SLOTS),
                      ("CPPRE_MAX_SCALARS", L.MAX_SCALARS), ("CPPRE_MAX_TAPS", L.MAX_TAPS)):
        assert int(re.search(rf"#define {name} (\d+)", header).group(1)) == val
    assert C.sizeof(L.Field) == 8 + 32 + 8      # d_data, stride[4], scale + offset (decode fold)
    assert C.sizeof(L.Program) == 12 + 4 + 72 + 72 + 48 + 48     # kind..n_slots (+pad), ksize, pad, taps, scalars
    assert C.sizeof(L.BC) == 48


def test_error_codes_map_to_reference_exception_types():
    lib = L.load()
    prog, epi = L.Program(), L.Epilogue()
    prog.kind = 99
    shape = (C.c_int64 * 4)(1, 1, 1, 1)
    fl = (L.Field * 1)()
    rc = lib.cppre_residual(C.byref(prog), shape, fl, C.byref(epi), 0, None)
    assert rc == L.ERR_BAD_ARG and b"unknown program kind" in lib.cppre_last_error()
    with pytest.raises(ValueError):
        L.check(rc)
    bc = ops.make_bc([7, 0, 0, 0], [1, 1, 1, 1], [0, 0, 0, 0])
    with pytest.raises(ValueError):
        ops.bc_padded_shape(bc, 4, 4)


def test_constructors_match_reference_kernels_bit_exact():
    g = load_golden("kernels")
    mods = _mods()
    n = 0
    for key in g.files:
        if key.startswith("err|"):
            continue
        module, dom, order, ty, scale = key.split("|")
        dom, order, ty, scale = ast.literal_eval(dom), ast.literal_eval(order), int(ty), float(scale)
        cls = mods[module].ConvOperator
        if module == "0d":
            op = cls(order, scale, ty)
        elif module == "spatial":
            op = cls(dom, order, scale, ty, device="cpu")
        else:
            op = cls(dom, order, scale, ty)
        want = g[key]
        if want.size == 0:
            assert not hasattr(op, "kernel"), key
        else:
            assert hasattr(op, "kernel"), key
            assert op.kernel.dtype == torch.float32
            np.testing.assert_array_equal(op.kernel.detach().numpy(), want, err_msg=key)
        n += 1
    assert n > 400


def test_constructor_error_and_attribute_behaviour():
    m = _mods()
    for mod in m.values():
        with pytest.raises(ValueError, match="Unknown Convolution Method"):
            (mod.ConvOperator(1, conv="bogus") if mod is m["0d"] else mod.ConvOperator("x", 1, conv="bogus"))
    D = m["2d"].ConvOperator("t", 2, scale=3.0)
    assert (D.domain, D.dims, D.order, D.axis) == ("t", 1, 2, 2)
    assert D.conv == D.convolution and m["2d"].ConvOperator("t", 1, conv="spectral").conv.__name__ == "spectral_convolution"
    assert torch.equal(D.stencil, m["2d"].get_stencil(1, 2))
    assert torch.equal(m["2d"].ConvOperator("y", 1).kernel, m["2d"].ConvOperator("t", 1).kernel)   # shipped defect
    E = m["2d"].ConvOperator()                       # additive composition target
    assert not hasattr(E, "kernel")
    with pytest.raises(AttributeError):
        E(torch.zeros(1, 3, 3, 3))
    E.kernel = D.kernel - 0.25 * m["2d"].ConvOperator(("x", "y"), 2).kernel
    assert E.kernel.shape == (3, 3, 3)
    # Spatial: default device is cuda; on a CUDA-less host the swallowed .to() failure leaves
    # the UNSCALED cpu stencil
    if not torch.cuda.is_available():
        S = m["spatial"].ConvOperator("x", 1, scale=5.0)
        np.testing.assert_array_equal(S.kernel.numpy()[:, 1], np.float32([-0.5, 0, 0.5]))
    assert m["2d"].kernel_3d(torch.ones(5, 5), 0)[1].sum() == 25
    with pytest.raises(ValueError):
        m["2d"].kernel_3d(torch.ones(3, 3), 3)
    assert m["2d"].pad_kernel(torch.zeros(2, 6, 7, 8), torch.ones(3, 3, 3)).shape == (6, 8, 7)
    assert m["1d"].pad_kernel(torch.zeros(2, 6, 7), torch.ones(3, 3)).shape == (6, 7)


def test_no_cpu_fallback_without_cuda():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    D = _mods()["2d"].ConvOperator("t", 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        D(torch.zeros(1, 4, 4, 4))
    from Neural_PDE.UQ.inductive_cp import calibrate
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        calibrate(np.zeros(20, np.float32), 20, 0.1)


def test_field_validation_happens_before_any_device_work():
    D = _mods()["2d"].ConvOperator("t", 1)
    with pytest.raises(RuntimeError):
        D(torch.zeros(4, 4, 4))                      # wrong rank
    with pytest.raises(RuntimeError, match="Float"):
        D(torch.zeros(1, 4, 4, 4, dtype=torch.float64))
    D.kernel = D.kernel.double()
    with pytest.raises(RuntimeError, match="Float"):
        D(torch.zeros(1, 4, 4, 4))
    D1 = _mods()["1d"].ConvOperator("x", 1)
    with pytest.raises(RuntimeError, match="channels"):
        D1(torch.zeros(2, 3, 5, 5))


def test_boundary_index_map_equals_oracle_padding():
    """The library's host-side index map, applied with NumPy, must reproduce the reference's
    sequential padding for every (left,right)x(top,bottom) type pair (tests/golden/bc.npz)."""
    from boundary_conditions import BoundaryManager
    g = load_golden("bc")
    x4 = g["x4"]
    types = [str(t) for t in g["types"]]
    assert types == BoundaryManager.SUPPORTED_TYPES
    for a in types:
        for c in types:
            b = BoundaryManager(5)
            b.set_boundary_type("left", a, 0.5)
            b.set_boundary_type("right", c, -0.5)
            b.set_boundary_type("top", c, 2.0)
            b.set_boundary_type("bottom", a, -2.0)
            rows, rowv, cols, colv = b.index_map(*x4.shape[-2:])
            want = g[f"mix|{a}|{c}"]
            assert want.shape[-2:] == (len(rows), len(cols)), (a, c)
            got = x4[..., np.maximum(rows, 0)[:, None], np.maximum(cols, 0)[None, :]].copy()
            # Dirichlet cells: the row constant wins over the column constant (rows padded last)
            got[..., :, cols < 0] = colv[cols < 0]
            got[..., rows < 0, :] = rowv[rows < 0, None]
            np.testing.assert_array_equal(got, want, err_msg=f"{a}|{c}")
    with pytest.raises(ValueError, match="Unknown side"):
        BoundaryManager(3).set_boundary_type("front", "periodic")
    with pytest.raises(ValueError, match="Unsupported boundary type"):
        BoundaryManager(3).set_boundary_type("left", "robin")
    assert str(g["err|side"]).startswith("Unknown side") and str(g["err|type"]).startswith("Unsupported")
    with pytest.raises(ValueError):                   # reflect needs pad < extent, like torch
        bm = BoundaryManager(7)
        bm.set_all_boundaries("symmetric")
        bm.index_map(3, 3)


def test_tma_eligibility_rules():
    """Which requests the dispatcher sends to the TMA star kernel (host-side classification)."""
    lib = L.load()

    def eligible(kernels, shape, strides=None, ptr=256, kind=L.PROG_LINEAR, pads=None, mode=L.EPI_SCORE):
        prog, epi = L.Program(), L.Epilogue()
        prog.kind, prog.n_fields, prog.n_slots = kind, L.PROGRAM_FIELDS[kind], len(kernels)
        keep = []
        for s, k in enumerate(kernels):
            k = np.ascontiguousarray(k, np.float32)
            keep.append(k)
            for d in range(3):
                prog.ksize[s][d] = k.shape[d]
                prog.pad[s][d] = (k.shape[d] // 2) if pads is None else pads[d]
            prog.h_taps[s] = k.ctypes.data
        fl = (L.Field * prog.n_fields)()
        B, N2, N1, N0 = shape
        st = strides or (N2 * N1 * N0, N1 * N0, N0, 1)
        for f in range(prog.n_fields):
            fl[f].d_data = ptr
            for d in range(4):
                fl[f].stride[d] = st[d]
        epi.mode, epi.flags, epi.d_scores, epi.d_out = mode, L.FLAG_ABS, 256, 256
        for d, v in enumerate((N2 * N1 * N0, N1 * N0, N0, 1)):
            epi.out_stride[d] = v
        return lib.cppre_residual_tma_eligible(C.byref(prog), (C.c_int64 * 4)(*shape), fl, C.byref(epi))

    wave = O.build_kernel("2d", "t", 2) - np.float32(0.25) * O.build_kernel("2d", ("x", "y"), 2)
    assert eligible([wave], (4, 8, 16, 128)) == 1
    assert eligible([wave], (4, 8, 16, 130)) == 0                  # rows not 16-byte multiples
    assert eligible([wave], (4, 8, 16, 128), ptr=260) == 0         # misaligned base
    assert eligible([wave], (4, 8, 16, 128), strides=(8 * 16 * 128, 1, 128 * 8, 8)) == 0   # t-innermost, raw
    dense = np.ones((3, 3, 3), np.float32)
    assert eligible([dense], (4, 8, 16, 128)) == 0                 # off-star taps -> generic kernel
    assert eligible([O.build_kernel("2d", ("x", "y"), 2, 1.0, 4)], (4, 8, 16, 128)) == 0    # 5^3
    k1 = O.build_kernel("1d", "t", 1).reshape(3, 1, 3)
    assert eligible([k1], (64, 10, 1, 1024)) == 1                  # 1-D+time: batch-rows mode
    assert eligible([wave], (4, 8, 16, 128), pads=(0, 0, 0)) == 0  # valid correlation
    ops2 = O.ops_2d()
    ks = [ops2[n] for n in ("D_t", "D_x", "D_y", "D_xx_yy")]
    assert eligible(ks, (2, 6, 16, 256), kind=L.PROG_NS_MOMENTUM, mode=L.EPI_STORE) == 1
    ks[3] = O.build_kernel("2d", ("x", "y"), 2, 1.0, 6)
    assert eligible(ks, (2, 6, 16, 256), kind=L.PROG_NS_MOMENTUM, mode=L.EPI_STORE) == 0


def test_calibrate_rank_and_level_match_numpy_formula():
    from Neural_PDE.UQ.inductive_cp import calibrate_level, calibrate_rank
    for n in (19, 20, 100, 256, 500, 4096):
        for alpha in np.arange(0.05, 0.95 + 0.1, 0.1):
            assert calibrate_level(n, alpha) == O.calibrate_quantile_level(n, alpha)
            try:
                want = O.calibrate_rank(n, n, alpha)
            except ValueError:
                with pytest.raises(ValueError, match="Quantiles must be in the range"):
                    calibrate_rank(n, n, alpha)
                continue
            assert calibrate_rank(n, n, alpha) == want
    # n != rows (the scripts pass len(other_array)): NumPy evaluates the index in the dtype of the score
    # array; both dtypes are reproduced (rows=14, n=13, alpha=0.5 -> 8 in float32, 7 in float64)
    assert calibrate_rank(14, 13, 0.5, np.float32) == 8 and calibrate_rank(14, 13, 0.5, np.float64) == 7
    rng = np.random.default_rng(0)
    for rows in list(range(2, 40)) + [100, 255, 256, 257, 1000]:
        for n in {rows, rows - 1, rows + 1, rows + 7, 2 * rows}:
            for alpha in np.arange(0.05, 0.96, 0.05):
                q = calibrate_level(n, alpha) if n >= 1 else 2.0
                if not 0 <= q <= 1:
                    continue
                for dt in (np.float32, np.float64):
                    x = rng.permutation(rows).astype(dt)
                    assert int(np.quantile(x, q, method="higher")) == calibrate_rank(rows, n, alpha, dt), (rows, n, alpha, dt)


def test_axis_canonicalisation_of_permuted_views():
    t = torch.zeros(2, 7, 8, 6).permute(0, 3, 1, 2)        # logical [BS,Nt,Nx,Ny], t innermost
    assert ops._axis_order(t) == [1, 2, 0]
    assert ops._axis_order(torch.zeros(2, 6, 7, 8)) == [0, 1, 2]
    assert ops._axis_order(torch.zeros(2, 6, 1, 8)) == [0, 1, 2]
    assert ops._axis_order(torch.zeros(2, 8, 6).permute(0, 2, 1).unsqueeze(2)) == [2, 1, 0]


def test_vector_ops_strict_reference_flag_raises_like_reference():
    import Utils.VectorConvOps as V
    g = load_golden("vector")
    assert V.Divergence().grad_x.kernel.shape == (3, 3, 3)
    V.STRICT_REFERENCE = True
    try:
        for cls in ("Divergence", "Gradient", "Curl", "Laplace"):
            with pytest.raises(ValueError, match="Unknown Convolution Method"):
                getattr(V, cls)()
            assert str(g[f"v2|{cls}|raises"]) == "ValueError: Unknown Convolution Method"
    finally:
        V.STRICT_REFERENCE = False
    a, b = torch.randn(2, 3, 4, 5), torch.randn(2, 3, 4, 5)
    assert torch.equal(V.cross((a, b), (b, b)), a * b + b * b)
    assert V.vectorize(a, b).shape == (2, 2, 3, 4, 5)


def test_integration_guide_covers_every_exported_entry_point():
    """INTEGRATION.md shows (or explains) the reference-side binding of every symbol include/cppre.h declares."""
    header = open(os.path.join(ROOT, "include", "cppre.h")).read()
    guide = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    names = set(re.findall(r"\b(cppre_[a-z0-9_]+)\s*\(", header))
    assert len(names) >= 25
    assert not [n for n in sorted(names) if n not in guide]


def test_decode_fold_host_mapping():
    """Normaliser-decode fold, host side: (scale, offset) forms, the MinMax-object form (decode(x) = (x - b) / a),
    per-variable entries routed to the program's fields by their index in ``vars`` (MHD induction reads variables
    1, 2, 4, 5), the coefficient field of the JOREK programs never decoded, malformed input refused."""
    import torch
    from cp_pre_b200 import programs
    mods = _mods()
    D = [mods["2d"].ConvOperator("t", 1), mods["2d"].ConvOperator("x", 1), mods["2d"].ConvOperator("y", 1)]

    class MinMax:
        a, b = torch.tensor([2.0, 4.0]), torch.tensor([1.0, -2.0])

    sc, of = programs.ResidualProgram._affine(MinMax())
    assert sc == [0.5, 0.25] and of == [-0.5, 0.5]
    assert programs.ResidualProgram._affine((3.0, -1.0)) == ([3.0], [-1.0])
    ind = programs.mhd("induction", *D)
    v = torch.zeros(2, 6, 4, 4, 4)
    sc6, of6 = [1.0, 2.0, 3.0, 4.0, 5.0, 6.0], [0.1, 0.2, 0.3, 0.4, 0.5, 0.6]
    assert ind._decode_per_field((sc6, of6), v, 4) == ([2.0, 3.0, 5.0, 6.0], [0.2, 0.3, 0.5, 0.6])
    assert ind._decode_per_field((2.0, 0.5), v, 4) == ([2.0] * 4, [0.5] * 4)          # scalars: every variable
    assert ind._decode_per_field(None, v, 4) is None
    with pytest.raises(ValueError):
        ind._decode_per_field(([1.0, 2.0], [0.0, 0.0]), v, 4)                          # 2 entries, 6 variables
    jor = programs.jorek("continuity", *D, D[1], D[2], torch.linspace(1, 2, 4), 3.4)
    s, o = jor._decode_per_field(([2.0, 3.0, 4.0], [0.0, 1.0, 2.0]), torch.zeros(2, 3, 4, 4, 4), 4)
    assert s == [2.0, 3.0, 4.0, 1.0] and o == [0.0, 1.0, 2.0, 0.0]                     # R is not decoded
    wave = programs.linear(D[0])
    assert wave._decode_per_field((0.25, 2.0), torch.zeros(2, 4, 4, 4), 1) == ([0.25], [2.0])
