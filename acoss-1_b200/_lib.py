# -*- coding: utf-8 -*-
"""
ctypes binding of csrc/libacoss_b200.so (C-ABI: include/acoss_b200.h).

There is NO CPU fallback: if the shared library is missing or no CUDA device is
usable, every compute entry point raises.  (`load_library()` itself works without
a GPU so that the CPU-only test-suite can check the exported symbols.)
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libacoss_b200.so")

c_f32p = ctypes.POINTER(ctypes.c_float)
c_f64p = ctypes.POINTER(ctypes.c_double)
c_i32p = ctypes.POINTER(ctypes.c_int32)
c_i64p = ctypes.POINTER(ctypes.c_int64)
c_u8p = ctypes.POINTER(ctypes.c_uint8)
c_vp = ctypes.c_void_p


class CrpParams(ctypes.Structure):
    _fields_ = [("m", ctypes.c_int32), ("tau", ctypes.c_int32), ("kappa", ctypes.c_float),
                ("oti", ctypes.c_int32), ("dis_onset", ctypes.c_float), ("dis_extension", ctypes.c_float),
                ("crp_mode", ctypes.c_int32)]


class PairStats(ctypes.Structure):
    _fields_ = [("ms_total", ctypes.c_double), ("ms_prep", ctypes.c_double), ("ms_crp", ctypes.c_double),
                ("ms_dp", ctypes.c_double), ("dp_cells", ctypes.c_int64), ("crp_cells", ctypes.c_int64),
                ("crp_bytes", ctypes.c_int64), ("n_launches", ctypes.c_int64), ("n_dp_launches", ctypes.c_int64),
                ("n_batches", ctypes.c_int64), ("n_exact_cells", ctypes.c_int64), ("n_fallback_rows", ctypes.c_int64),
                ("crp_kernel_mask", ctypes.c_int64), ("dp_kernel_mask", ctypes.c_int64)]

    CRP_KERNELS = {1: "exact", 2: "band", 4: "band_tc", 8: "gemm", 16: "split"}
    DP_KERNELS = {1: "rowpar", 2: "generic"}

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["crp_kernels"] = [n for b, n in self.CRP_KERNELS.items() if d["crp_kernel_mask"] & b]
        d["dp_kernels"] = [n for b, n in self.DP_KERNELS.items() if d["dp_kernel_mask"] & b]
        return d


# every symbol include/acoss_b200.h declares: (restype, argtypes)
SYMBOLS = {
    "acoss_ctx_create": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(c_vp)]),
    "acoss_ctx_destroy": (None, [c_vp]),
    "acoss_last_error": (ctypes.c_char_p, [c_vp]),
    "acoss_version": (ctypes.c_char_p, []),
    "acoss_ctx_stream": (c_vp, [c_vp]),
    "acoss_ctx_synchronize": (ctypes.c_int, [c_vp]),
    "acoss_ctx_set_stream": (ctypes.c_int, [c_vp, c_vp]),
    "acoss_ctx_set_option": (ctypes.c_int, [c_vp, ctypes.c_char_p, ctypes.c_int64]),
    "acoss_upload_tracks": (ctypes.c_int, [c_vp, c_f32p, c_i64p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32]),
    "acoss_median_downsample": (ctypes.c_int, [c_vp, c_f32p, c_i64p, ctypes.c_int32, ctypes.c_int32, c_f32p, c_i64p]),
    "acoss_serra09_pairs": (ctypes.c_int, [c_vp, c_i32p, ctypes.c_int64, ctypes.POINTER(CrpParams), c_f32p, c_f32p]),
    "acoss_serra09_pairs_dev": (ctypes.c_int, [c_vp, c_i32p, ctypes.c_int64, ctypes.POINTER(CrpParams), c_vp, c_vp]),
    "acoss_get_pair_stats": (ctypes.c_int, [c_vp, ctypes.POINTER(PairStats)]),
    "acoss_debug_counters": (ctypes.c_int, [c_vp, ctypes.POINTER(ctypes.c_uint64)]),
    "acoss_debug_mma_probe": (ctypes.c_int, [c_vp, ctypes.c_int32, c_i64p]),
    "acoss_debug_counters_ext": (ctypes.c_int, [c_vp, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int32]),
    "acoss_crp_pair": (ctypes.c_int, [c_vp, ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(CrpParams), c_u8p,
                                      c_i32p, c_i32p, c_f32p, c_f32p, c_i32p]),
    "acoss_cover_similarity_from_crp": (ctypes.c_int, [c_vp, c_u8p, c_i32p, c_i32p, ctypes.c_int32, ctypes.c_int32,
                                                       ctypes.c_float, ctypes.c_float, c_f32p]),
    "acoss_cover_similarity_from_crp_noalloc": (ctypes.c_int, [c_vp, c_vp, c_i32p, c_i32p, ctypes.c_int32, ctypes.c_int32, c_vp]),
    "acoss_scores_to_matrix_dev": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, c_vp]),
    "acoss_normalize_by_length_dev": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_vp, ctypes.c_int32]),
    "acoss_eval_statistics": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_i32p, c_i32p, c_i32p, c_i32p,
                                             ctypes.c_int32, c_f64p, c_f64p]),
    "acoss_np_pairwise_sum": (ctypes.c_double, [c_f64p, ctypes.c_int64]),
    "acoss_sw_constrained_from_binary": (ctypes.c_int, [c_vp, c_u8p, c_i32p, c_i32p, ctypes.c_int32, c_f64p]),
    "acoss_snf": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                 ctypes.c_double, c_f64p, c_vp]),
    "acoss_simple_upload": (ctypes.c_int, [c_vp, c_f64p, c_i64p, ctypes.c_int32]),
    "acoss_simple_preprocess": (ctypes.c_int, [c_vp, c_f32p, c_i64p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                               ctypes.POINTER(ctypes.c_double), ctypes.c_int32, ctypes.POINTER(ctypes.c_double), c_i64p]),
    "acoss_simple_pairs": (ctypes.c_int, [c_vp, c_i32p, ctypes.c_int64, ctypes.c_int32, c_f64p]),
    "acoss_ftm2d_allpairs": (ctypes.c_int, [c_vp, c_f64p, ctypes.c_int32, ctypes.c_int32, c_f64p]),
    "acoss_ftm2d_shingles": (ctypes.c_int, [c_vp, c_f32p, c_i64p, c_i64p, c_i64p, ctypes.c_int32, ctypes.c_double, ctypes.c_int32,
                                            ctypes.c_double, c_f64p]),
    "acoss_serra09_topk": (ctypes.c_int, [c_vp, c_i32p, ctypes.c_int32, c_i32p, ctypes.c_int32, ctypes.POINTER(CrpParams), ctypes.c_int32,
                                          ctypes.c_int32, ctypes.c_int32, c_i32p, c_f32p]),
    "acoss_snf_ex": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                    ctypes.c_double, c_f64p, c_vp, c_f64p, c_f64p]),
    "acoss_csm": (ctypes.c_int, [c_vp, c_f32p, ctypes.c_int32, c_f32p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, c_f32p]),
    "acoss_oti": (ctypes.c_int, [c_vp, c_f32p, c_f32p, ctypes.c_int32, c_i32p]),
    "acoss_csm_to_binary": (ctypes.c_int, [c_vp, c_f32p, ctypes.c_int32, ctypes.c_int32, ctypes.c_double, c_u8p]),
    "acoss_wcsm": (ctypes.c_int, [c_vp, c_f32p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, c_f32p]),
    "acoss_earlyfusion_block_features": (ctypes.c_int, [c_vp, c_f32p, ctypes.c_int32, c_i64p, c_f32p, c_i64p, c_i64p, c_i64p, ctypes.c_int32,
                                                        ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, c_f32p, c_f32p, c_f32p, c_f32p, c_i64p]),
    "acoss_earlyfusion_upload": (ctypes.c_int, [c_vp, c_f32p, ctypes.c_int32, c_f32p, ctypes.c_int32, c_f32p, ctypes.c_int32,
                                                c_f32p, c_i64p, ctypes.c_int32]),
    "acoss_earlyfusion_pairs": (ctypes.c_int, [c_vp, c_i32p, ctypes.c_int64, ctypes.c_double, ctypes.c_int32, c_f64p]),
    "acoss_earlyfusion_pair_matrices": (ctypes.c_int, [c_vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_double, ctypes.c_int32,
                                                       c_f32p, c_u8p, c_f64p]),
    "acoss_earlyfusion_stats": (ctypes.c_int, [c_vp, c_f64p, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)]),
    "acoss_dev_alloc": (ctypes.c_int, [c_vp, ctypes.c_int64, ctypes.POINTER(c_vp)]),
    "acoss_dev_free": (ctypes.c_int, [c_vp, c_vp]),
    "acoss_memcpy_h2d": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int64]),
    "acoss_memcpy_d2h": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int64]),
}

_LIB = None


class AcossError(RuntimeError):
    """Raised for any non-zero return code of the C-ABI (code in .code)."""

    def __init__(self, code, msg):
        RuntimeError.__init__(self, "acoss_b200 error %d: %s" % (code, msg))
        self.code = code


def load_library():
    """dlopen the CUDA library and bind every declared symbol.  Raises if absent."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("acoss_b200: %s is missing; run __graft_entry__.build() "
                              "(nvcc -gencode arch=compute_100a,code=sm_100a).  There is no CPU fallback." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def _np(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def _ptr(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct)) if a is not None else None


class Context(object):
    """One context per GPU (acoss_ctx).  All methods raise AcossError on failure."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = c_vp()
        rc = self.lib.acoss_ctx_create(int(device), ctypes.byref(h))
        if rc != 0:
            raise AcossError(rc, (self.lib.acoss_last_error(None) or b"").decode())
        self.h = h
        self.device = int(device)
        self.lengths = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.acoss_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise AcossError(rc, (self.lib.acoss_last_error(self.h) or b"").decode())

    def synchronize(self):
        self._ck(self.lib.acoss_ctx_synchronize(self.h))

    def set_option(self, key, value):
        """acoss_ctx_set_option: "crp_budget_bytes", "inject_alloc_failures" (include/acoss_b200.h)."""
        self._ck(self.lib.acoss_ctx_set_option(self.h, key.encode(), int(value)))

    # ---- tracks ---------------------------------------------------------------
    @staticmethod
    def pack_tracks(tracks):
        lens = np.array([t.shape[0] for t in tracks], dtype=np.int64)
        offs = np.zeros(len(tracks) + 1, np.int64)
        offs[1:] = np.cumsum(lens)
        frames = np.empty((int(offs[-1]), 12), np.float32)
        for t, o in zip(tracks, offs[:-1]):
            frames[o:o + t.shape[0]] = t
        return frames, offs

    def upload_tracks(self, tracks=None, m=9, tau=1, frames=None, offsets=None):
        if frames is None:
            frames, offsets = self.pack_tracks(tracks)
        frames = _np(frames, np.float32)
        offsets = _np(offsets, np.int64)
        self._ck(self.lib.acoss_upload_tracks(self.h, _ptr(frames, ctypes.c_float), _ptr(offsets, ctypes.c_int64),
                                              len(offsets) - 1, int(m), int(tau)))
        self.lengths = np.diff(offsets).astype(np.int32)
        self.h2d_bytes = frames.nbytes + offsets.nbytes

    def median_downsample(self, raw_tracks, fac):
        frames, offs = self.pack_tracks([_np(t, np.float32) for t in raw_tracks])
        n = len(raw_tracks)
        out_offs = np.zeros(n + 1, np.int64)
        total = int(sum((t.shape[0] + fac - 1) // fac for t in raw_tracks))
        out = np.empty((total, 12), np.float32)
        self._ck(self.lib.acoss_median_downsample(self.h, _ptr(frames, ctypes.c_float), _ptr(offs, ctypes.c_int64), n, int(fac),
                                                  _ptr(out, ctypes.c_float), _ptr(out_offs, ctypes.c_int64)))
        return [out[out_offs[i]:out_offs[i + 1]] for i in range(n)]

    # ---- pairs ------------------------------------------------------------------
    @staticmethod
    def params(m=9, tau=1, kappa=0.095, oti=True, dis_onset=0.5, dis_extension=0.5, crp_mode=0):
        return CrpParams(int(m), int(tau), float(kappa), int(bool(oti)), float(dis_onset), float(dis_extension), int(crp_mode))

    def serra09_pairs(self, pairs, params=None, want_dmax=False, want_qmax=True):
        """Host in / host out.  Returns (qmax|None, dmax|None)."""
        params = params or self.params()
        pairs = _np(pairs, np.int32).reshape(-1, 2)
        P = pairs.shape[0]
        q = np.zeros(P, np.float32) if want_qmax else None
        d = np.zeros(P, np.float32) if want_dmax else None
        self._ck(self.lib.acoss_serra09_pairs(self.h, _ptr(pairs, ctypes.c_int32), P, ctypes.byref(params),
                                              _ptr(q, ctypes.c_float), _ptr(d, ctypes.c_float)))
        return q, d

    def serra09_pairs_dev(self, pairs, params, out_q_ptr, out_d_ptr):
        """Host pairs in, device pointers out (ints; 0/None = not wanted).  Asynchronous."""
        pairs = _np(pairs, np.int32).reshape(-1, 2)
        self._ck(self.lib.acoss_serra09_pairs_dev(self.h, _ptr(pairs, ctypes.c_int32), pairs.shape[0], ctypes.byref(params),
                                                  c_vp(out_q_ptr or 0), c_vp(out_d_ptr or 0)))

    def set_stream(self, cuda_stream):
        """Order every later call on the caller's stream (an int cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream);
        None / 0 returns to the context's own stream."""
        self._ck(self.lib.acoss_ctx_set_stream(self.h, c_vp(cuda_stream or 0)))

    def serra09_topk(self, queries, db, k, params=None, which="qmax", norm="serra09"):
        """Per-query top-k of queries x db (track indices of the uploaded store) without the Q x N matrix: returns
        (idx (Q, k) int32 positions in `db`, -1 = none; score (Q, k) float32).  which: 'qmax' | 'dmax';
        norm: None (raw, higher first) | 'serra09' (score / sqrt(len), higher first) | 'chen' (sqrt(len) / score, lower first)."""
        params = params or self.params()
        queries = _np(queries, np.int32).ravel(); db = _np(db, np.int32).ravel()
        idx = np.zeros((len(queries), int(k)), np.int32)
        sc = np.zeros((len(queries), int(k)), np.float32)
        self._ck(self.lib.acoss_serra09_topk(self.h, _ptr(queries, ctypes.c_int32), len(queries), _ptr(db, ctypes.c_int32), len(db),
                                             ctypes.byref(params), {"qmax": 0, "dmax": 1}[which], {None: -1, "serra09": 0, "chen": 1}[norm],
                                             int(k), _ptr(idx, ctypes.c_int32), _ptr(sc, ctypes.c_float)))
        return idx, sc

    def pair_stats(self):
        s = PairStats()
        self._ck(self.lib.acoss_get_pair_stats(self.h, ctypes.byref(s)))
        return s.as_dict()

    def debug_counters(self):
        out = (ctypes.c_uint64 * 16)()
        self._ck(self.lib.acoss_debug_counters(self.h, out))
        return [int(x) for x in out]

    def debug_counters_ext(self, n=48):
        out = (ctypes.c_uint64 * n)()
        self._ck(self.lib.acoss_debug_counters_ext(self.h, out, n))
        return [int(x) for x in out]

    def crp_pair(self, q, r, params=None):
        """(crp uint8 (M,N), thr_q, thr_r, oti) of one uploaded pair (test hook)."""
        params = params or self.params()
        nq, nr = int(self.lengths[q]), int(self.lengths[r])
        buf = np.zeros(max(nq, 1) * max(nr, 1), np.uint8)
        tq = np.zeros(max(nq, 1), np.float32)
        tr = np.zeros(max(nr, 1), np.float32)
        M, N, o = ctypes.c_int32(0), ctypes.c_int32(0), ctypes.c_int32(0)
        self._ck(self.lib.acoss_crp_pair(self.h, int(q), int(r), ctypes.byref(params), _ptr(buf, ctypes.c_uint8),
                                         ctypes.byref(M), ctypes.byref(N), _ptr(tq, ctypes.c_float), _ptr(tr, ctypes.c_float),
                                         ctypes.byref(o)))
        return buf[:M.value * N.value].reshape(M.value, N.value).copy(), tq[:M.value].copy(), tr[:N.value].copy(), o.value

    @staticmethod
    def _pack_mats(mats):
        mats = [_np(x, np.uint8) for x in mats]
        Ms = np.array([x.shape[0] for x in mats], np.int32)
        Ns = np.array([x.shape[1] for x in mats], np.int32)
        flat = np.concatenate([x.ravel() for x in mats]) if mats else np.zeros(0, np.uint8)
        return _np(flat, np.uint8), Ms, Ns

    def cover_similarity_from_crp(self, crps, alignment="serra09", dis_onset=0.5, dis_extension=0.5):
        flat, Ms, Ns = self._pack_mats(crps)
        out = np.zeros(len(crps), np.float32)
        a = {"serra09": 0, "chen17": 1}[alignment]
        self._ck(self.lib.acoss_cover_similarity_from_crp(self.h, _ptr(flat, ctypes.c_uint8), _ptr(Ms, ctypes.c_int32),
                                                          _ptr(Ns, ctypes.c_int32), len(crps), a, dis_onset, dis_extension,
                                                          _ptr(out, ctypes.c_float)))
        return out

    def cover_similarity_from_crp_noalloc(self, crps_dev_ptr, Ms, Ns, out_dev_ptr, alignment="serra09"):
        """Allocation-free Qmax / Dmax on device-resident 0/1 matrices (include/acoss_b200.h): device pointers as ints."""
        Ms, Ns = _np(Ms, np.int32), _np(Ns, np.int32)
        a = {"serra09": 0, "chen17": 1}[alignment]
        self._ck(self.lib.acoss_cover_similarity_from_crp_noalloc(self.h, c_vp(crps_dev_ptr), _ptr(Ms, ctypes.c_int32),
                                                                  _ptr(Ns, ctypes.c_int32), len(Ms), a, c_vp(out_dev_ptr)))

    def sw_constrained(self, mats):
        flat, Ms, Ns = self._pack_mats(mats)
        out = np.zeros(len(mats), np.float64)
        self._ck(self.lib.acoss_sw_constrained_from_binary(self.h, _ptr(flat, ctypes.c_uint8), _ptr(Ms, ctypes.c_int32),
                                                           _ptr(Ns, ctypes.c_int32), len(mats), _ptr(out, ctypes.c_double)))
        return out

    # ---- N x N stage (device pointers as ints) -----------------------------------
    def scores_to_matrix_dev(self, pairs_ptr, scores_ptr, n_pairs, n, symmetric, D_ptr):
        self._ck(self.lib.acoss_scores_to_matrix_dev(self.h, c_vp(pairs_ptr), c_vp(scores_ptr), int(n_pairs), int(n),
                                                     int(bool(symmetric)), c_vp(D_ptr)))

    def normalize_by_length_dev(self, D_ptr, n, lengths_ptr, mode):
        self._ck(self.lib.acoss_normalize_by_length_dev(self.h, c_vp(D_ptr), int(n), c_vp(lengths_ptr), int(mode)))

    def eval_statistics(self, D_ptr, n, order, csize, cstart, tops):
        order, csize, cstart, tops = (_np(x, np.int32) for x in (order, csize, cstart, tops))
        out = np.zeros(4 + len(tops), np.float64)
        ranks = np.zeros(n, np.float64)
        self._ck(self.lib.acoss_eval_statistics(self.h, c_vp(D_ptr), int(n), _ptr(order, ctypes.c_int32), _ptr(csize, ctypes.c_int32),
                                                _ptr(cstart, ctypes.c_int32), _ptr(tops, ctypes.c_int32), len(tops),
                                                _ptr(out, ctypes.c_double), _ptr(ranks, ctypes.c_double)))
        return out, ranks

    # ---- similarity network fusion -------------------------------------------------
    def snf(self, mats, K=5, niters=5, reg_diag=1, want_Ws=True):
        """doSimilarityFusion on a list of (n, n) distance matrices (float32 or float64, all the same
        dtype).  Returns (Ws list | None, fused float64 (n, n))."""
        dt = np.float64 if any(np.asarray(m).dtype == np.float64 for m in mats) else np.float32
        stack = np.ascontiguousarray(np.stack([np.asarray(m, dtype=dt) for m in mats]))
        nm, n = stack.shape[0], stack.shape[1]
        fused = np.zeros((n, n), np.float64)
        Ws = np.zeros_like(stack) if want_Ws else None
        self._ck(self.lib.acoss_snf(self.h, stack.ctypes.data_as(c_vp), 0 if dt == np.float32 else 1, nm, n, int(K), int(niters),
                                    float(reg_diag), _ptr(fused, ctypes.c_double), Ws.ctypes.data_as(c_vp) if want_Ws else None))
        return ([Ws[k] for k in range(nm)] if want_Ws else None), fused

    def snf_stages(self, mats, K, niters=0, reg_diag=1, affinities=False, want=("fused",)):
        """acoss_snf_ex: the stages of doSimilarityFusion on a list of (n, n) matrices (distances, or affinities when
        `affinities`).  `want` names the outputs: 'fused', 'Ws', 'P', 'S' (S dense).  Returns a dict."""
        dt = np.float64 if any(np.asarray(m).dtype == np.float64 for m in mats) else np.float32
        stack = np.ascontiguousarray(np.stack([np.asarray(m, dtype=dt) for m in mats]))
        nm, n = stack.shape[0], stack.shape[1]
        if stack.ndim != 3 or stack.shape[2] != n:
            raise ValueError("square matrices expected")
        out = {}
        if "fused" in want: out["fused"] = np.zeros((n, n), np.float64)
        if "Ws" in want: out["Ws"] = np.zeros_like(stack)
        if "P" in want: out["P"] = np.zeros((nm, n, n), np.float64)
        if "S" in want: out["S"] = np.zeros((nm, n, n), np.float64)
        self._ck(self.lib.acoss_snf_ex(self.h, stack.ctypes.data_as(c_vp), 0 if dt == np.float32 else 1, 1 if affinities else 0, nm, n,
                                       int(K), int(niters), float(reg_diag), _ptr(out.get("fused"), ctypes.c_double),
                                       out["Ws"].ctypes.data_as(c_vp) if "Ws" in out else None,
                                       _ptr(out.get("P"), ctypes.c_double), _ptr(out.get("S"), ctypes.c_double)))
        return out

    # ---- free functions of acoss/algorithms/utils ------------------------------------------
    def csm(self, X, Y, cosine=False, roll=0):
        """get_csm / get_csm_cosine (roll: get_csm_blocked_oti's shift of X) -> (M, N) float32."""
        X = _np(X, np.float32); Y = _np(Y, np.float32)
        if X.ndim != 2 or Y.ndim != 2 or X.shape[1] != Y.shape[1]:
            raise ValueError("X (M, d) and Y (N, d) expected")
        out = np.zeros((X.shape[0], Y.shape[0]), np.float32)
        self._ck(self.lib.acoss_csm(self.h, _ptr(X, ctypes.c_float), X.shape[0], _ptr(Y, ctypes.c_float), Y.shape[0], X.shape[1],
                                    1 if cosine else 0, int(roll), _ptr(out, ctypes.c_float)))
        return out

    def oti(self, C1, C2):
        """get_oti for (n, 12) float32 pairs of chroma vectors -> int32 (n,)."""
        C1 = _np(C1, np.float32).reshape(-1, 12); C2 = _np(C2, np.float32).reshape(-1, 12)
        out = np.zeros(len(C1), np.int32)
        self._ck(self.lib.acoss_oti(self.h, _ptr(C1, ctypes.c_float), _ptr(C2, ctypes.c_float), len(C1), _ptr(out, ctypes.c_int32)))
        return out

    def csm_to_binary(self, D, kappa):
        D = _np(D, np.float32)
        out = np.zeros(D.shape, np.uint8)
        self._ck(self.lib.acoss_csm_to_binary(self.h, _ptr(D, ctypes.c_float), D.shape[0], D.shape[1], float(kappa), _ptr(out, ctypes.c_uint8)))
        return out

    def wcsm(self, C, k1, k2):
        C = _np(C, np.float32)
        out = np.zeros(C.shape, np.float32)
        self._ck(self.lib.acoss_wcsm(self.h, _ptr(C, ctypes.c_float), C.shape[0], C.shape[1], int(k1), int(k2), _ptr(out, ctypes.c_float)))
        return out

    # ---- SiMPle / FTM2D ---------------------------------------------------------------
    def simple_preprocess(self, raw, win=200, skip=100, win_len_smooth=4):
        """Simple.load_features + Simple.smooth for a list of raw (T_t, 12) float32 chroma arrays, on the device
        (simple_silva.py:34-43, 56-66).  Returns a list of (12, int(T_t / skip)) float64 arrays."""
        raw = [_np(r, np.float32).reshape(-1, 12) for r in raw]
        offs = np.zeros(len(raw) + 1, np.int64)
        offs[1:] = np.cumsum([r.shape[0] for r in raw])
        flat = _np(np.concatenate(raw, 0) if raw else np.zeros((0, 12)), np.float32)
        k = np.arange(win_len_smooth + 2)
        taps = 0.5 - 0.5 * np.cos(2.0 * np.pi * k / (win_len_smooth + 1))          # symmetric Hann (fftbins=False)
        taps = _np(taps / np.sum(taps), np.float64)
        cols = np.array([r.shape[0] // skip for r in raw], np.int64)
        out = np.zeros(12 * max(int(cols.sum()), 1), np.float64)
        co = np.zeros(len(raw) + 1, np.int64)
        self._ck(self.lib.acoss_simple_preprocess(self.h, _ptr(flat, ctypes.c_float), _ptr(offs, ctypes.c_int64), len(raw), int(win), int(skip),
                                                  _ptr(taps, ctypes.c_double), len(taps), _ptr(out, ctypes.c_double), _ptr(co, ctypes.c_int64)))
        return [out[12 * co[t]:12 * co[t + 1]].reshape(12, -1).copy() for t in range(len(raw))]

    def simple_upload(self, feats):
        """feats: list of (12, L_t) float64 arrays (Simple.load_features output)."""
        feats = [_np(f, np.float64) for f in feats]
        offs = np.zeros(len(feats) + 1, np.int64)
        offs[1:] = np.cumsum([f.shape[1] for f in feats])
        flat = np.concatenate([f.ravel() for f in feats]) if feats else np.zeros(0)
        flat = _np(flat, np.float64)
        self._ck(self.lib.acoss_simple_upload(self.h, _ptr(flat, ctypes.c_double), _ptr(offs, ctypes.c_int64), len(feats)))

    def simple_pairs(self, pairs, sslen=10):
        pairs = _np(pairs, np.int32).reshape(-1, 2)
        out = np.zeros(pairs.shape[0], np.float64)
        self._ck(self.lib.acoss_simple_pairs(self.h, _ptr(pairs, ctypes.c_int32), pairs.shape[0], int(sslen), _ptr(out, ctypes.c_double)))
        return out

    def ftm2d_allpairs(self, shingles):
        S = _np(shingles, np.float64)
        n, d = S.shape
        D = np.zeros((n, n), np.float64)
        self._ck(self.lib.acoss_ftm2d_allpairs(self.h, _ptr(S, ctypes.c_double), n, d, _ptr(D, ctypes.c_double)))
        return D

    def ftm2d_shingles(self, raw, onsets, pwr=1.96, win=75, C=5):
        """FTM2D.load_features for a list of raw (T_t, 12) float32 chroma arrays and their beat onsets (frame indices), on the
        device (ftm2d.py:52-66).  Returns (n_tracks, 12 * win) float64 unit shingles."""
        raw = [_np(r, np.float32).reshape(-1, 12) for r in raw]
        offs = np.zeros(len(raw) + 1, np.int64)
        offs[1:] = np.cumsum([r.shape[0] for r in raw])
        flat = _np(np.concatenate(raw, 0), np.float32)
        ons = [_np(np.asarray(o).ravel(), np.int64) for o in onsets]
        ooff = np.zeros(len(raw) + 1, np.int64)
        ooff[1:] = np.cumsum([len(o) for o in ons])
        oflat = _np(np.concatenate(ons) if ons else np.zeros(0), np.int64)
        if len(oflat) == 0:
            oflat = np.zeros(1, np.int64)
        out = np.zeros((len(raw), 12 * int(win)), np.float64)
        self._ck(self.lib.acoss_ftm2d_shingles(self.h, _ptr(flat, ctypes.c_float), _ptr(offs, ctypes.c_int64), _ptr(oflat, ctypes.c_int64),
                                               _ptr(ooff, ctypes.c_int64), len(raw), float(pwr), int(win), float(C), _ptr(out, ctypes.c_double)))
        return out

    # ---- EarlyFusionTralie ---------------------------------------------------------------
    def earlyfusion_block_features(self, mfccs, chromas, onsets, blocksize=20, mfccs_per_block=50, chromas_per_block=40):
        """EarlyFusion.load_features on the device for a list of tracks: mfccs[t] (T, n_mfcc) float32 (= feats['mfcc_htk'].T),
        chromas[t] (T, 12) float32, onsets[t] beat frame indices.  Returns a list of dicts 'mfccs', 'ssms', 'chromas',
        'chroma_med' (earlyfusion_traile.py:104-150)."""
        n = len(mfccs)
        mf = [_np(m, np.float32) for m in mfccs]
        n_mfcc = int(mf[0].shape[1])
        ch = [_np(c, np.float32).reshape(-1, 12) for c in chromas]
        ons = [_np(np.asarray(o).ravel(), np.int64) for o in onsets]

        def offs(lens):
            o = np.zeros(n + 1, np.int64)
            o[1:] = np.cumsum(lens)
            return o
        moff, coff, ooff = offs([m.shape[0] for m in mf]), offs([c.shape[0] for c in ch]), offs([len(o) for o in ons])
        mflat, cflat, oflat = _np(np.concatenate(mf, 0), np.float32), _np(np.concatenate(ch, 0), np.float32), _np(np.concatenate(ons), np.int64)
        boff = np.zeros(n + 1, np.int64)
        args = [self.h, _ptr(mflat, ctypes.c_float), n_mfcc, _ptr(moff, ctypes.c_int64), _ptr(cflat, ctypes.c_float), _ptr(coff, ctypes.c_int64),
                _ptr(oflat, ctypes.c_int64), _ptr(ooff, ctypes.c_int64), n, int(blocksize), int(mfccs_per_block), int(chromas_per_block)]
        self._ck(self.lib.acoss_earlyfusion_block_features(*(args + [None, None, None, None, _ptr(boff, ctypes.c_int64)])))
        NB = int(boff[-1])
        om = np.zeros((NB, mfccs_per_block * n_mfcc), np.float32)
        os_ = np.zeros((NB, mfccs_per_block * (mfccs_per_block - 1) // 2), np.float32)
        oc = np.zeros((NB, chromas_per_block * 12), np.float32)
        omed = np.zeros((n, 12), np.float32)
        self._ck(self.lib.acoss_earlyfusion_block_features(*(args + [_ptr(om, ctypes.c_float), _ptr(os_, ctypes.c_float), _ptr(oc, ctypes.c_float),
                                                                     _ptr(omed, ctypes.c_float), _ptr(boff, ctypes.c_int64)])))
        return [{"mfccs": om[boff[t]:boff[t + 1]].copy(), "ssms": os_[boff[t]:boff[t + 1]].copy(), "chromas": oc[boff[t]:boff[t + 1]].copy(),
                 "chroma_med": omed[t].copy()} for t in range(n)]

    def earlyfusion_upload(self, block_feats):
        """block_feats: list of dicts with 'mfccs', 'ssms', 'chromas' (n_blocks, d) and 'chroma_med' (12)."""
        nb = [int(f['mfccs'].shape[0]) for f in block_feats]
        offs = np.zeros(len(nb) + 1, np.int64)
        offs[1:] = np.cumsum(nb)
        cat = {k: _np(np.concatenate([np.asarray(f[k], np.float32) for f in block_feats]), np.float32) for k in ('mfccs', 'ssms', 'chromas')}
        cmed = _np(np.stack([np.asarray(f['chroma_med'], np.float32) for f in block_feats]), np.float32)
        self.ef_blocks = np.array(nb, np.int64)
        self._ck(self.lib.acoss_earlyfusion_upload(self.h, _ptr(cat['mfccs'], ctypes.c_float), cat['mfccs'].shape[1],
                                                   _ptr(cat['ssms'], ctypes.c_float), cat['ssms'].shape[1],
                                                   _ptr(cat['chromas'], ctypes.c_float), cat['chromas'].shape[1],
                                                   _ptr(cmed, ctypes.c_float), _ptr(offs, ctypes.c_int64), len(nb)))

    def earlyfusion_pairs(self, pairs, kappa=0.1, K=10):
        """(P, 4) float64 scores: mfccs, ssms, chromas, early."""
        pairs = _np(pairs, np.int32).reshape(-1, 2)
        out = np.zeros((pairs.shape[0], 4), np.float64)
        self._ck(self.lib.acoss_earlyfusion_pairs(self.h, _ptr(pairs, ctypes.c_int32), pairs.shape[0], float(kappa), int(K),
                                                  _ptr(out, ctypes.c_double)))
        return out

    def earlyfusion_pair_matrices(self, i, j, kappa=0.1, K=10):
        M, N = int(self.ef_blocks[i]), int(self.ef_blocks[j])
        csm = np.zeros((4, M, N), np.float32)
        bn = np.zeros((4, M, N), np.uint8)
        sc = np.zeros(4, np.float64)
        self._ck(self.lib.acoss_earlyfusion_pair_matrices(self.h, int(i), int(j), float(kappa), int(K), _ptr(csm, ctypes.c_float),
                                                          _ptr(bn, ctypes.c_uint8), _ptr(sc, ctypes.c_double)))
        return csm, bn, sc

    def earlyfusion_stats(self):
        ms = np.zeros(3, np.float64)
        fl = ctypes.c_double(0)
        cells = ctypes.c_int64(0)
        self._ck(self.lib.acoss_earlyfusion_stats(self.h, _ptr(ms, ctypes.c_double), ctypes.byref(fl), ctypes.byref(cells)))
        return {"ms_csm": ms[0], "ms_select": ms[1], "ms_sw": ms[2], "gemm_flop": fl.value, "cells": cells.value}
