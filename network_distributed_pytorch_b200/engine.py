"""CUDA engine: owns plans (device tables) and issues the per-phase kernels through the C ABI.

The reducer's host logic talks to an engine through a handful of phase calls on torch
tensors; this is the only engine the package ships and it needs a CUDA device and the built
library.  (The test-suite injects a CPU test double built on the oracle to exercise the host
logic and the gloo collectives without a GPU; that double lives under tests/, not here.)
"""
import ctypes

import torch

from . import _lib


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptrs(tensors):
    return _lib.ptr_array([t.data_ptr() for t in tensors])


class CudaPlan:
    """Packed layout + tile tables for one fixed list of gradient shapes (reducer.py:69-98)."""

    def __init__(self, lib, device, mat_dims, rank, vec_lens):
        self._lib = lib
        self.device = device
        self.n_mat = len(mat_dims)
        self.n_vec = len(vec_lens)
        handle = ctypes.c_void_p()
        n = _lib.i64_array([d[0] for d in mat_dims])
        m = _lib.i64_array([d[1] for d in mat_dims])
        v = _lib.i64_array(list(vec_lens))
        _lib.check(lib.psgd_plan_create(self.n_mat, n, m, int(rank), self.n_vec, v,
                                        -1 if device is None else device.index, ctypes.byref(handle)), lib)
        self._h = handle
        self.p_elems = lib.psgd_p_elems(handle)
        self.q_elems = lib.psgd_q_elems(handle)
        self.vec_elems = lib.psgd_vec_elems(handle)
        self.workspace_bytes = lib.psgd_workspace_bytes(handle)
        self.mats = []
        buf = (ctypes.c_int64 * 8)()
        for i in range(self.n_mat):
            _lib.check(lib.psgd_matrix_info(handle, i, buf), lib)
            self.mats.append(dict(n=buf[0], m=buf[1], r=buf[2], p_off=buf[3], q_off=buf[4], ws_off=buf[5],
                                  row_blocks=buf[6], cols_per_tile=buf[7]))
        self._workspace = None

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.psgd_plan_destroy(h)
            except Exception:
                pass

    # -- host-only inspection -------------------------------------------------------------
    def tiles(self, phase):
        out = []
        buf = (ctypes.c_int32 * 6)()
        for i in range(self._lib.psgd_tile_count(self._h, phase)):
            _lib.check(self._lib.psgd_tile_get(self._h, phase, i, buf), self._lib)
            out.append(tuple(buf))
        return out

    # -- per step ---------------------------------------------------------------------------
    def set_ptrs(self, mat_in, mat_out, mat_mem, vec_in, vec_out, grad=None, err=None, vec_grad=None, vec_err=None):
        if vec_grad is not None:  # fused error-feedback add: operands for matrices and vectors
            arrays = (_ptrs(mat_in), _ptrs(mat_out), _ptrs(mat_mem), _ptrs(vec_in), _ptrs(vec_out),
                      _ptrs(grad), _ptrs(err), _ptrs(vec_grad), _ptrs(vec_err))
            _lib.check(self._lib.psgd_set_ptrs_fused(self._h, *arrays, _stream(self.device)), self._lib)
            return
        arrays = (_ptrs(mat_in), _ptrs(mat_out), _ptrs(mat_mem), _ptrs(vec_in), _ptrs(vec_out),
                  _ptrs(grad) if grad is not None else None, _ptrs(err) if err is not None else None)
        _lib.check(self._lib.psgd_set_ptrs(self._h, *arrays, _stream(self.device)), self._lib)

    def workspace(self):
        if self._workspace is None:
            self._workspace = torch.empty(self.workspace_bytes, dtype=torch.uint8, device=self.device)
        return self._workspace

    def pack_vec(self, flat):
        _lib.check(self._lib.psgd_pack_vec(self._h, flat.data_ptr(), _stream(self.device)), self._lib)

    def pack_vec_fused(self, flat):
        _lib.check(self._lib.psgd_pack_vec_fused(self._h, flat.data_ptr(), _stream(self.device)), self._lib)

    def p_from_m(self, p_mem, q_mem, fuse_add=False, q_div=1.0):
        _lib.check(self._lib.psgd_p_from_m(self._h, p_mem.data_ptr(), q_mem.data_ptr(), int(fuse_add),
                                           float(q_div), _stream(self.device)), self._lib)

    def p_from_m_pack(self, p_mem, q_mem, flat, fuse_add=False, q_div=1.0):
        """p_from_m with the flat-buffer gather folded into the same launch."""
        _lib.check(self._lib.psgd_p_from_m_pack(self._h, p_mem.data_ptr(), q_mem.data_ptr(), int(fuse_add),
                                                float(q_div), flat.data_ptr(), _stream(self.device)), self._lib)

    def p_from_m_orth(self, p_mem, q_mem, flat=None, fuse_add=False, q_div=1.0):
        """Single worker: P = M Q and its Gram-Schmidt in one call (no all-reduce sits between them); the q_from_mt
        call that follows overlaps what is left of the orthogonalisation."""
        _lib.check(self._lib.psgd_p_from_m_orth(self._h, p_mem.data_ptr(), q_mem.data_ptr(), int(fuse_add), float(q_div),
                                                flat.data_ptr() if flat is not None else None, _stream(self.device)),
                   self._lib)

    # -- N > 1 inside one NVLink domain: the all-reduces ride inside these launches (peer.PeerExchange, fused modes) --
    def p_from_m_peer(self, peer, p_mem, q_mem, fuse_add=False):
        _lib.check(self._lib.psgd_p_from_m_peer(self._h, peer.handle, p_mem.data_ptr(), q_mem.data_ptr(), int(fuse_add),
                                                _stream(self.device)), self._lib)

    def p_from_m_push(self, peer, q_mem, fuse_add=False, with_flat=True, reduce_q_first=False):
        _lib.check(self._lib.psgd_p_from_m_push(self._h, peer.handle, q_mem.data_ptr(), int(fuse_add), int(with_flat),
                                                int(reduce_q_first), _stream(self.device)), self._lib)

    def q_from_mt_push(self, peer, p_mem):
        _lib.check(self._lib.psgd_q_from_mt_push(self._h, peer.handle, p_mem.data_ptr(), self.workspace().data_ptr(),
                                                 _stream(self.device)), self._lib)

    def orthogonalize_peer(self, peer, p_mem, flat_out):
        _lib.check(self._lib.psgd_orthogonalize_peer(self._h, peer.handle, p_mem.data_ptr(),
                                                     flat_out.data_ptr() if flat_out is not None else None,
                                                     _stream(self.device)), self._lib)

    def finalize_peer(self, peer, p_mem, q_mem, flat):
        _lib.check(self._lib.psgd_finalize_peer(self._h, peer.handle, p_mem.data_ptr(), q_mem.data_ptr(), flat.data_ptr(),
                                                _stream(self.device)), self._lib)

    def finalize_unpack(self, p_mem, q_mem, n_workers, flat):
        """finalize with the scatter of the flat buffer folded into the same launch."""
        _lib.check(self._lib.psgd_finalize_unpack(self._h, p_mem.data_ptr(), q_mem.data_ptr(), float(n_workers),
                                                  flat.data_ptr(), _stream(self.device)), self._lib)

    def orthogonalize(self, p_mem, n_workers):
        _lib.check(self._lib.psgd_orthogonalize(self._h, p_mem.data_ptr(), float(n_workers),
                                                _stream(self.device)), self._lib)

    def q_from_mt(self, p_mem, q_mem):
        _lib.check(self._lib.psgd_q_from_mt(self._h, p_mem.data_ptr(), q_mem.data_ptr(),
                                            self.workspace().data_ptr(), _stream(self.device)), self._lib)

    def finalize(self, p_mem, q_mem, n_workers):
        _lib.check(self._lib.psgd_finalize(self._h, p_mem.data_ptr(), q_mem.data_ptr(), float(n_workers),
                                           _stream(self.device)), self._lib)

    def finish(self, flat, q_mem, n_workers):
        _lib.check(self._lib.psgd_finish(self._h, flat.data_ptr() if flat is not None else None,
                                         q_mem.data_ptr() if q_mem is not None else None, float(n_workers),
                                         _stream(self.device)), self._lib)

    # -- timing -----------------------------------------------------------------------------
    def timing_enable(self, phases=True):
        """phases: True = all, False = off, or an iterable of phase names (see _lib.PHASE_NAMES)."""
        if phases is True:
            mask = (1 << _lib.N_PHASES) - 1
        elif not phases:
            mask = 0
        else:
            mask = sum(1 << _lib.PHASE_NAMES.index(p) for p in phases)
        _lib.check(self._lib.psgd_timing_enable(self._h, mask), self._lib)

    def timing_read(self):
        ms = (ctypes.c_double * _lib.N_PHASES)()
        cnt = (ctypes.c_int64 * _lib.N_PHASES)()
        _lib.check(self._lib.psgd_timing_read(self._h, ms, cnt), self._lib)
        return {name: (ms[i], cnt[i]) for i, name in enumerate(_lib.PHASE_NAMES)}


class CudaEngine:
    """The product engine.  Construction fails loudly when CUDA or the library is missing."""

    name = "cuda-sm100a"

    def __init__(self, device):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError(
                f"PowerSGDReducer (B200 build) needs a CUDA device, got '{device}'. "
                "There is no CPU fallback; the CPU restatement under oracle/ is test infrastructure only.")
        if not torch.cuda.is_available():
            raise RuntimeError("CUDA is not available: the B200 PowerSGD reducer cannot run on this machine")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = device
        self.lib = _lib.load()

    def make_plan(self, mat_dims, rank, vec_lens):
        return CudaPlan(self.lib, self.device, mat_dims, rank, vec_lens)

    def empty(self, n):
        return torch.empty(n, device=self.device, dtype=torch.float32)

    def orthogonalize_matrix(self, matrix):
        n, r = matrix.shape
        _lib.check(self.lib.psgd_orthogonalize_one(matrix.data_ptr(), n, r, self.device.index,
                                                   _stream(self.device)), self.lib)

    def launch_count(self):
        return self.lib.psgd_launch_count()


def host_plan(mat_dims, rank, vec_lens):
    """Layout-only plan (no CUDA call): packed offsets and tile tables, for inspection/tests."""
    return CudaPlan(_lib.load(), None, mat_dims, rank, vec_lens)
