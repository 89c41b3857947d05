"""The C-ABI library without a GPU: it loads, exports every symbol include/psgd_b200.h declares,
and its host-only plan reproduces the upstream packed layout and tiles every matrix exactly."""
import ctypes
import os
import re

import numpy as np
import pytest

from network_distributed_pytorch_b200 import _lib, workloads
from network_distributed_pytorch_b200.build import build
from network_distributed_pytorch_b200.engine import host_plan

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build()
    return _lib.load()


def header_symbols():
    text = open(os.path.join(ROOT, "include", "psgd_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(psgd_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in psgd_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes table and header disagree"
    assert lib.psgd_abi_version() == _lib.ABI_VERSION


def test_compute_entry_points_fail_loudly_on_host_only_plan(lib):
    plan = host_plan([(8, 6)], 2, [3])
    rc = lib.psgd_p_from_m(plan._h, None, None, 0, ctypes.c_float(1.0), None)
    assert rc != 0 and b"host-only" in lib.psgd_last_error()
    h = ctypes.c_void_p()
    assert lib.psgd_plan_create(1, _lib.i64_array([4]), _lib.i64_array([4]), 0, 0, _lib.i64_array([]), -1,
                                ctypes.byref(h)) != 0  # rank 0 rejected
    assert lib.psgd_plan_create(1, _lib.i64_array([64]), _lib.i64_array([64]), 33, 0, _lib.i64_array([]), -1,
                                ctypes.byref(h)) != 0  # beyond PSGD_MAX_RANK
    assert b"PSGD_MAX_RANK" in lib.psgd_last_error()


def test_peer_allreduce_entry_points_validate_before_touching_the_gpu(lib):
    one = _lib.ptr_array([16])
    for fn, extra, tail in ((lib.psgd_peer_allreduce, (), ()),
                            (lib.psgd_peer_allreduce_mc, (ctypes.c_void_p(16),), (ctypes.c_float(1.0), 0, ctypes.c_void_p(16)))):
        args = lambda world, rank, n=8, out=16: ((one, one) + extra + (ctypes.c_void_p(out), n, rank, world, 256, 1) + tail
                                                 + (0, None))
        assert fn(*args(9, 0)) != 0 and b"world size" in lib.psgd_last_error()
        assert fn(*args(1, 1)) != 0 and b"rank" in lib.psgd_last_error()
        assert fn(*args(1, 0, n=-1)) != 0
        assert fn(*args(1, 0, n=0)) == 0  # nothing to do, nothing launched
    assert lib.psgd_peer_allreduce_mc(one, one, None, ctypes.c_void_p(16), 8, 0, 1, 256, 1, ctypes.c_float(1.0), 0,
                                      ctypes.c_void_p(16), 0, None) != 0
    assert lib.psgd_peer_allreduce_mc(one, one, ctypes.c_void_p(16), ctypes.c_void_p(16), 8, 0, 1, 256, 1,
                                      ctypes.c_float(0.0), 0, ctypes.c_void_p(16), 0, None) != 0 and b"out_div" in lib.psgd_last_error()
    assert lib.psgd_peer_allreduce_mc(one, one, ctypes.c_void_p(16), ctypes.c_void_p(16), 8, 0, 1, 256, 1,
                                      ctypes.c_float(1.0), 0, None, 0, None) != 0 and b"blocks_done" in lib.psgd_last_error()


def test_peer_handles_validate_before_touching_the_gpu(lib):
    """psgd_peer_create / psgd_peer_create_push: geometry errors are reported, nothing is launched."""
    one = _lib.ptr_array([16, 32])
    h = ctypes.c_void_p()
    mk = lambda world, rank, mode, mc=16: lib.psgd_peer_create(world, rank, 0, one, one, ctypes.c_void_p(mc), one, one,
                                                              ctypes.c_void_p(mc), 256, mode, ctypes.byref(h))
    assert mk(1, 0, 2) != 0 and b"world size" in lib.psgd_last_error()
    assert mk(2, 2, 2) != 0 and b"rank" in lib.psgd_last_error()
    assert mk(2, 0, 7) != 0 and b"mode" in lib.psgd_last_error()
    assert mk(2, 0, 1, mc=0) != 0 and b"multicast" in lib.psgd_last_error()
    push = lambda stride, mc=16: lib.psgd_peer_create_push(2, 0, 0, ctypes.c_void_p(16), ctypes.c_void_p(mc), stride, one,
                                                           ctypes.c_void_p(16), ctypes.c_void_p(16), 8, one, 256,
                                                           ctypes.byref(h))
    assert push(6) != 0 and b"stride" in lib.psgd_last_error()
    assert push(8, mc=0) != 0 and b"multicast" in lib.psgd_last_error()
    assert lib.psgd_orthogonalize_peer(None, None, None, None, None) != 0
    assert lib.psgd_finalize_peer(None, None, None, None, None, None) != 0
    assert lib.psgd_p_from_m_push(None, None, None, 0, 1, 0, None) != 0
    assert lib.psgd_peer_destroy(None) == 0


def split(shapes):
    mats = [(s[0], int(np.prod(s[1:]))) for s in shapes if len(s) > 1]
    vecs = [int(np.prod(s)) if len(s) else 1 for s in shapes if len(s) <= 1]
    return mats, vecs


@pytest.mark.parametrize("name,rank", [("resnet18", 2), ("distilbert", 4), ("distilbert", 1), ("resnet50", 4),
                                       ("bert-large", 8), ("resnet152", 4), ("distilbert", 16), ("distilbert", 3)])
def test_packed_layout_and_tiles(lib, name, rank):
    shapes = workloads.WORKLOADS[name]()
    mats, vecs = split(shapes)
    st = workloads.stats(shapes, rank)
    plan = host_plan(mats, rank, vecs)
    assert plan.p_elems == st["e_p"] and plan.q_elems == st["e_q"] and plan.vec_elems == st["e_1"]
    po = qo = 0
    for (n, m), info in zip(mats, plan.mats):  # reducer.py:86-98 running offsets, no padding
        r = min(n, m, rank)
        assert (info["n"], info["m"], info["r"], info["p_off"], info["q_off"]) == (n, m, r, po, qo)
        po += n * r
        qo += m * r
    # row tiles of K1 / K4 partition the rows of every matrix exactly once
    for phase in (_lib.PHASE_P_FROM_M, _lib.PHASE_FINALIZE):
        cover = [0] * len(mats)
        for mat, row0, nrows, _, ncols, _ in plan.tiles(phase):
            assert row0 == cover[mat] and nrows > 0 and ncols == mats[mat][1]
            cover[mat] += nrows
        assert cover == [n for n, _ in mats]
    # K3 tiles partition (rows x columns); workspace slices do not overlap
    seen = {}
    for mat, row0, nrows, col0, ncols, blk in plan.tiles(_lib.PHASE_Q_FROM_MT):
        assert nrows > 0 and ncols > 0
        seen.setdefault(mat, []).append((row0, nrows, col0, ncols, blk))
    ws_end = 0
    for i, ((n, m), info) in enumerate(zip(mats, plan.mats)):
        area = sum(t[1] * t[3] for t in seen[i])
        assert area == n * m
        assert max(t[4] for t in seen[i]) + 1 == info["row_blocks"]
        assert info["ws_off"] >= ws_end and info["ws_off"] % 4 == 0
        ws_end = info["ws_off"] + info["row_blocks"] * m * info["r"]
    assert plan.workspace_bytes >= 4 * ws_end


@pytest.mark.parametrize("name,rank", [("distilbert", 4), ("distilbert", 8), ("distilbert", 16), ("bert-large", 8),
                                       ("resnet50", 4), ("tall_mix", 4), ("tall_mix", 16)])
def test_gram_schmidt_job_table(lib, name, rank):
    """K2's plan-time routing: every matrix is orthogonalised by exactly one group of CTAs; groups are powers of two,
    start at a multiple of their size inside a hardware cluster (so they never straddle one), hold all rows of their
    matrix in registers (4 rows per thread) and no CTA of a group keeps more than 1024 rows unless the group already
    is a whole cluster; a matrix too tall for a whole cluster goes to the generic kernel."""
    if name == "tall_mix":
        mats = [(4096, 8), (4097, 8), (9000, 8), (30522, 8), (65536, 8), (65537, 8), (3, 8), (2049, 8)]
    else:
        mats, _ = split(workloads.WORKLOADS[name]())
    plan = host_plan(mats, rank, [])
    R = 1
    while R < min(rank, max(min(n, m) for n, m in mats)):
        R *= 2
    threads = 512 if R >= 16 else 1024
    jobs = plan.tiles(_lib.PHASE_ORTH)
    groups = {}
    for pos, (mat, cbase, gsize, grank, cluster, generic) in enumerate(jobs):
        groups.setdefault(mat, []).append((pos, cbase, gsize, grank, cluster, generic))
    assert sorted(groups) == [i for i, (n, m) in enumerate(mats) if n and m]
    for mat, ctas in groups.items():
        n = mats[mat][0]
        pos0, cbase, gsize, _, cluster, generic = ctas[0]
        assert len(ctas) == gsize and [c[3] for c in ctas] == list(range(gsize))
        assert all(c[1] == cbase and c[2] == gsize and c[4] == cluster and c[5] == generic for c in ctas)
        if generic:
            assert n > cluster * 2 * threads * 4 or R > 16  # nothing register-resident could have taken it
            continue
        assert [c[0] for c in ctas] == list(range(pos0, pos0 + gsize))  # consecutive CTAs of the launch
        assert gsize & (gsize - 1) == 0 and cbase % gsize == 0 and cbase + gsize <= cluster
        assert pos0 % cluster == cbase
        per_cta = -(-n // gsize)
        assert per_cta <= threads * 4
        assert per_cta <= 1024 or gsize == cluster or n <= threads * 4 and gsize == 1


def test_ragged_and_degenerate_shapes(lib):
    mats = [(1, 5), (5, 1), (64, 147), (3, 7), (1000, 12), (2, 4096 * 3)]
    plan = host_plan(mats, 4, [1, 7])
    assert [d["r"] for d in plan.mats] == [1, 1, 4, 3, 4, 2]
    for phase in (_lib.PHASE_P_FROM_M, _lib.PHASE_FINALIZE):
        rows = {}
        for mat, row0, nrows, *_ in plan.tiles(phase):
            rows[mat] = rows.get(mat, 0) + nrows
        assert rows == {i: n for i, (n, _) in enumerate(mats)}
    cols = {}
    for mat, row0, nrows, col0, ncols, blk in plan.tiles(_lib.PHASE_Q_FROM_MT):
        cols[mat] = cols.get(mat, 0) + nrows * ncols
    assert cols == {i: n * m for i, (n, m) in enumerate(mats)}
