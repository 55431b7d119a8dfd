"""CPU-side tests of the product's host logic (no GPU, no compute calls on the device):
the C-ABI library loads and exports every declared symbol, the host class builds and
decodes the compressed format bit-for-bit like the reference (golden dumps + oracle),
and the slab partitioner behaves."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import sparsematrix_b200 as sb
from sparsematrix_b200 import _capi, gen
from oracle import oracle as O
from helpers import assert_bits_equal, golden_names, load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    declared = set()
    for hdr in ("sblas_b200.h", "sblas_b200_host.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        declared |= set(re.findall(r"SBLAS_B200_API[^;(]*?\b(sblas_\w+)\s*\(", text))
    assert len(declared) >= 25
    L = C.CDLL(_capi.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in include/ but not exported"
    assert declared == set(_capi.SYMBOLS), "ctypes table and headers disagree"
    assert sb.lib().sblas_b200_abi_version() == 1


def test_library_exports_the_reference_free_functions_under_their_mangled_names():
    # kernel.h:42-62 / kernel.cc:802-811 of the reference: six explicit instantiations.  Code compiled against the
    # reference header must link against this build, so the names are the Itanium manglings of those instantiations.
    L = C.CDLL(sb.LIB_PATH)
    for name in ("_Z27sblas_beta_operation_kernelIfEvPT_iiiS0_",
                 "_Z18sblas_trans_kernelIfEvPT_iiiS1_i",
                 "_Z22sblas_kernel_operationIhhfLi8EEviiiPT1_iS1_iS0_PT_PT0_iS1_i",
                 "_Z28sblas_kernel_operation_naiveIhhfLi8EEviiiPT1_iS1_iS0_PT_PT0_iS1_i",
                 "_Z28sblas_kernel_operation_transIhhfLi8EEviiiPT1_iS1_iS0_PT_PT0_iS1_i",
                 "_Z31sblas_kernel_operation_trans_exIhhfLi8EEviiiPT1_iS1_iS0_PT_PT0_iS1_i"):
        assert hasattr(L, name), name
    # and the class itself, as the reference instantiates it (sparse-matrix.cc:315), in the static archive that
    # carries the reference's library name (src/sparse/Makefile:9)
    import shutil
    import subprocess
    archive = os.path.join(os.path.dirname(sb.LIB_PATH), "libsparsematrix.a")
    if shutil.which("nm") and os.path.exists(archive):
        syms = subprocess.run(["nm", "--defined-only", archive], capture_output=True, text=True).stdout
        for name in ("_ZN5sblas12SparseMatrixIhhfLi0ELi8EE9AddMatMatEPfiiS2_iff", "_ZN5sblas12SparseMatrixIhhfLi0ELi8EE8CopyFormEPKhiiiPKfiNS_15SBLAS_TRANSPOSEE",
                     "_ZN5sblas12SparseMatrixIhhfLi0ELi8EE6CopyToEPfiNS_15SBLAS_TRANSPOSEE", "_Z18sblas_trans_kernelIfEvPT_iiiS1_i"):
            assert name in syms, name


def test_product_does_not_reference_oracle():
    # the oracle is test infrastructure: nothing under sparsematrix_b200/ may import or link it
    for dirpath, _, files in os.walk(os.path.join(ROOT, "sparsematrix_b200")):
        for f in files:
            if f.endswith((".py", ".cc", ".cu", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "sblas_oracle" not in text and "from oracle" not in text, (dirpath, f)


@pytest.mark.parametrize("name", golden_names())
def test_host_class_matches_reference_dump(name):
    g = load_golden(name)
    S = sb.SparseMatrix()
    S.CopyForm(g["codes"], int(g["rows"]), int(g["cols"]), int(g["stride"]), g["vals"], int(g["T"]), int(g["trans"]))
    assert (S.NumRows(), S.NumCols()) == (int(g["s_rows"]), int(g["s_cols"]))
    pos, code, table, T, bb, bib = S.members()
    assert np.array_equal(pos, g["pos"]) and np.array_equal(code, g["code"])
    assert_bits_equal(table, g["table"], "val_table_")
    assert np.array_equal(bb, g["block_bounds"]) and np.array_equal(bib, g["block_index_bounds"])
    assert_bits_equal(S.CopyTo(int(g["stride_notrans"]), 0), g["dense_notrans"], "CopyTo NoTrans")
    assert_bits_equal(S.CopyTo(int(g["stride_trans"]), 1), g["dense_trans"], "CopyTo Trans")


@pytest.mark.parametrize("trans", [0, 1])
@pytest.mark.parametrize("shape", [(1023, 511, 512, 63, 0.25), (300, 700, 700, 255, 0.01), (5, 1, 1, 3, 0.9)])
def test_host_class_matches_oracle(shape, trans):
    rows, cols, stride, T, dens = shape
    codes = np.full((rows, stride), 255, np.uint8)
    codes[:, :cols] = gen.codes_bernoulli(rows, cols, dens, seed=rows + cols, T=T)
    tab = gen.codebook(T, 2)
    S = sb.SparseMatrix(codes, rows, cols, stride, tab, T, trans)
    s = O.copyform(codes, rows, cols, stride, tab, T, bool(trans))
    pos, code, table, T2, bb, bib = S.members()
    assert s.same_as(O.Stream(pos, code, table, T2, bb, bib, S.NumRows(), S.NumCols()))
    assert_bits_equal(S.CopyTo(stride, trans), O.copyto(s, stride, bool(trans)), "CopyTo")
    S2 = sb.SparseMatrix(codes, rows, cols, stride, tab, T, trans)
    assert S == S2
    S2.Destroy()
    assert (S2.NumRows(), S2.NumCols()) == (0, 0) and not (S == S2)


def test_empty_table_and_bad_table_size():
    S = sb.SparseMatrix(np.zeros(12, np.uint8), 3, 4, 4, np.zeros(1, np.float32), 0)
    assert (S.NumRows(), S.NumCols()) == (0, 0)          # sparse-matrix.cc:26
    with pytest.raises(ValueError):
        S.CopyForm(np.zeros(12, np.uint8), 3, 4, 4, np.zeros(300, np.float32), 256)


def test_partition_balances_entries_on_panel_boundaries():
    n, k = 2048, 300
    codes = gen.codes_bernoulli(n, k, 0.05, seed=5)
    codes[0:512] = 255                                      # two empty leading panels
    codes[1536:1792, ::2] = 255                             # one half-density panel
    S = sb.SparseMatrix(codes, n, k, k, gen.codebook(), 255, 1)
    _, _, _, _, bb, bib = S.members()
    for parts in (1, 2, 3, 4, 8, 16):
        sp = sb.partition(bb, bib, n, parts)
        assert sp[0] == 0 and sp[-1] == n and np.all(np.diff(sp) >= 0)
        assert all(int(x) % 256 == 0 for x in sp[1:-1])
    sp = sb.partition(bb, bib, n, 2)
    cnt = np.zeros(8, np.int64)
    for (r, c), (s, e) in zip(bb, bib):
        cnt[c // 256] = e - s
    left = cnt[: sp[1] // 256].sum()
    best = min(abs(cnt[:b].sum() * 2 - cnt.sum()) for b in range(9))
    assert abs(left * 2 - cnt.sum()) == best
    # cols not a multiple of 256
    sp = sb.partition(bb[:1], bib[:1], 700, 4)
    assert sp[-1] == 700 and np.all(np.diff(sp) >= 0)


def test_argument_errors_are_reported_without_a_device():
    L = sb.lib()
    out = C.c_void_p()
    rc = L.sblas_b200_upload(None, None, 5, None, 300, None, None, 0, 4, 4, 0, C.byref(out))
    assert rc == _capi.EINVAL and b"valid_table_size" in L.sblas_b200_last_error()
    rc = L.sblas_b200_addmatmat_host(None, None, 1, 1, None, 1, 1.0, 1.0)
    assert rc == _capi.EINVAL
    # peer-visible output: argument checks first, then "no device" -- never a host allocation instead
    handle = (C.c_uint8 * 64)()
    assert L.sblas_b200_peer_alloc(0, 0, C.byref(out), handle) == _capi.EINVAL
    assert L.sblas_b200_peer_alloc(0, 64, None, handle) == _capi.EINVAL
    assert L.sblas_b200_peer_close(C.c_void_p(1234)) == _capi.EINVAL
    assert L.sblas_b200_peer_free(C.c_void_p(1234)) == _capi.EINVAL
    # ranges the application mapped itself: bookkeeping only, overlaps refused, each call pairs with its own undo
    base = 0x7000_0000_0000
    assert L.sblas_b200_peer_register(None, 64) == _capi.EINVAL
    assert L.sblas_b200_peer_register(C.c_void_p(base), 0) == _capi.EINVAL
    assert L.sblas_b200_peer_register(C.c_void_p(base), 4096) == _capi.OK
    assert L.sblas_b200_peer_register(C.c_void_p(base), 4096) == _capi.EINVAL             # the same range again
    assert L.sblas_b200_peer_register(C.c_void_p(base + 4095), 16) == _capi.EINVAL        # tail overlap
    assert L.sblas_b200_peer_register(C.c_void_p(base - 8), 16) == _capi.EINVAL           # head overlap
    assert L.sblas_b200_peer_register(C.c_void_p(base + 4096), 16) == _capi.OK            # adjacent is fine
    assert L.sblas_b200_peer_close(C.c_void_p(base)) == _capi.EINVAL                      # not a _peer_open mapping
    assert L.sblas_b200_peer_unregister(C.c_void_p(base + 8)) == _capi.EINVAL             # not a range start
    assert L.sblas_b200_peer_unregister(C.c_void_p(base)) == _capi.OK
    assert L.sblas_b200_peer_unregister(C.c_void_p(base)) == _capi.EINVAL
    assert L.sblas_b200_peer_unregister(C.c_void_p(base + 4096)) == _capi.OK
    import torch
    if not torch.cuda.is_available():
        assert L.sblas_b200_peer_alloc(0, 64, C.byref(out), handle) == _capi.ECUDA
        assert L.sblas_b200_host_alloc(64, C.byref(out)) == _capi.ECUDA      # no malloc in disguise
    assert L.sblas_b200_host_alloc(0, C.byref(out)) == _capi.EINVAL
    assert L.sblas_b200_host_free(None) == _capi.OK


def test_save_and_load_round_trip_and_reject_corruption(tmp_path):
    """Dump / load of the compressed members (sblas_host_save / _load): equal object after a round trip,
    nothing changes when the file is damaged, truncated or missing."""
    from sparsematrix_b200 import gen
    n, k = 300, 700
    codes = gen.codes_bernoulli(n, k, 0.03, seed=71)
    codes[:, 256:512] = 255                                    # an empty panel (Trans: S columns are the n rows; keep a gap in r instead)
    tab = gen.codebook()
    for trans in (sb.SparseMatrix.NoTrans, sb.SparseMatrix.Trans):
        S = sb.SparseMatrix(codes, n, k, k, tab, 255, trans)
        path = str(tmp_path / f"m{trans}.sblas")
        S.Save(path)
        L = sb.SparseMatrix()
        L.Load(path)
        assert L == S and L.NumRows() == S.NumRows() and L.NumCols() == S.NumCols()
        for x, y in zip(L.members(), S.members()):
            assert np.array_equal(np.asarray(x), np.asarray(y))
        raw = bytearray(open(path, "rb").read())
        keep = sb.SparseMatrix(codes[:10], 10, k, k, tab, 255, trans)
        before = keep.members()
        for name, data in (("flipped", bytes(raw[:100]) + bytes([raw[100] ^ 0x40]) + bytes(raw[101:])),
                           ("truncated", bytes(raw[:-9])), ("trailing", bytes(raw) + b"x"), ("magic", b"NOTSBLAS" + bytes(raw[8:]))):
            bad = str(tmp_path / f"{name}.sblas")
            open(bad, "wb").write(data)
            with pytest.raises(sb.SblasError) as ei:
                keep.Load(bad)
            assert ei.value.code == _capi.EFORMAT, name
            for x, y in zip(keep.members(), before):
                assert np.array_equal(np.asarray(x), np.asarray(y)), name
        with pytest.raises(sb.SblasError) as ei:
            keep.Load(str(tmp_path / "missing.sblas"))
        assert ei.value.code == _capi.EIO
    # a crafted file whose checksum matches but whose positions run past the matrix, or repeat a position: rejected before
    # anything can decode it (the host CopyTo has no bounds check, like the reference's)
    def fnv1a(data):
        h = 14695981039346656037
        for byte in data:
            h = ((h ^ byte) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        return h
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, sb.SparseMatrix.NoTrans)
    path = str(tmp_path / "good.sblas")
    S.Save(path)
    raw = bytearray(open(path, "rb").read())
    pos, code, table, T, bb, bib = S.members()
    pos_off = 32 + 4 * (T + 1) + 16 * bb.shape[0]
    assert bytes(raw[pos_off: pos_off + 16]) == pos[:16].tobytes()
    for name, edit in (("past the last row", lambda b: b.__setitem__(slice(pos_off + int(bib[-1, 0]), pos_off + int(bib[-1, 1])), b"\xff" * int(bib[-1, 1] - bib[-1, 0]))),
                       ("repeated position", lambda b: b.__setitem__(pos_off + 1, 0))):
        body = bytearray(raw[:-8])
        edit(body)
        crafted = str(tmp_path / "crafted.sblas")
        open(crafted, "wb").write(bytes(body) + fnv1a(body).to_bytes(8, "little"))
        with pytest.raises(sb.SblasError) as ei:
            keep.Load(crafted)
        assert ei.value.code == _capi.EFORMAT, name
    E = sb.SparseMatrix()                                      # empty object
    E.Save(str(tmp_path / "empty.sblas"))
    F = sb.SparseMatrix(codes, n, k, k, tab, 255, 0)
    F.Load(str(tmp_path / "empty.sblas"))
    assert F == E and F.NumRows() == 0 and F.NumCols() == 0


def test_partition_csr_balances_nnz_on_aligned_rows():
    # the CSR analogue of sblas_b200_partition (config 4 cannot go through the panel format): nnz-balanced, aligned splits
    n = 100000
    lens = gen.powerlaw_row_lengths(n, seed=5, x_min=8, cap=1 << 14)
    row_ptr = np.concatenate(([0], np.cumsum(lens))).astype(np.int32)
    for parts in (1, 2, 3, 8):
        sp = sb.partition_csr(row_ptr, parts, 256)
        assert sp[0] == 0 and sp[-1] == n and np.all(np.diff(sp) >= 0)
        assert np.all(sp[1:-1] % 256 == 0)
        share = np.diff(row_ptr[sp].astype(np.int64))
        assert share.sum() == row_ptr[-1]
        assert share.max() <= row_ptr[-1] / parts + 256 * lens.max()          # within one aligned step of the ideal share
    # everything in one row: the other slabs are empty, never negative
    rp = np.zeros(1001, np.int32); rp[501:] = 7777
    sp = sb.partition_csr(rp, 4, 256)
    assert sp[0] == 0 and sp[-1] == 1000 and np.all(np.diff(sp) >= 0)
    with pytest.raises(sb.SblasError):
        sb.partition_csr(row_ptr, 2, 100)                                     # alignment must be a power of two
