"""ctypes binding of oracle/_build/liboracle.so (oracle.c, the plain-C CPU restatement).

TEST INFRASTRUCTURE ONLY.  Same contracts as the numpy functions of oracle/oracle.py; used where
the numpy version is too slow (np.add.at over millions of edges) - full-size property checks and
bench.py's CPU legs.  Build: `make -C oracle oracle`.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
_MSG = {"copy_src": 0, "src_mul_edge": 1, "copy_edge": 2}

_LIB = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(LIB_PATH)
    return _LIB


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


_I = ctypes.c_int64


def in_degrees(dst, n):
    dst = _i64(dst)
    out = np.empty(n, dtype=np.int64)
    lib().oracle_in_degrees(_I(n), _I(len(dst)), _p(dst), _p(out))
    return out


def get_adj_csr(src, dst, n, transpose=False):
    src, dst = _i64(src), _i64(dst)
    key, other = (src, dst) if transpose else (dst, src)
    e = len(src)
    indptr = np.empty(n + 1, dtype=np.int64)
    indices = np.empty(e, dtype=np.int64)
    eid = np.empty(e, dtype=np.int64)
    lib().oracle_get_adj_csr(_I(n), _I(e), _p(key), _p(other), _p(indptr), _p(indices), _p(eid))
    return indptr, indices, eid


def _prep(msg_op, src, x, w):
    e = len(src)
    if msg_op == "copy_edge":
        w2 = _f32(w).reshape(e, -1)
        return None, w2, w2.shape[1], w2.shape[1]
    x2 = _f32(x).reshape(len(x), -1)
    if msg_op == "copy_src":
        return x2, None, x2.shape[1], 1
    w2 = _f32(w).reshape(e, -1)
    return x2, w2, x2.shape[1], w2.shape[1]


def spmm_sum(src, dst, n_out, msg_op, x=None, w=None):
    """(n_out, feat) float32; w of width w_width broadcasts over trailing groups of feat/w_width."""
    src, dst = _i64(src), _i64(dst)
    x2, w2, feat, ww = _prep(msg_op, src, x, w)
    out = np.empty((n_out, feat), dtype=np.float32)
    lib().oracle_spmm_sum(_I(n_out), _I(len(src)), _p(src), _p(dst), ctypes.c_int(_MSG[msg_op]),
                          _p(x2), _p(w2), _I(feat), _I(ww), _p(out))
    return out


def spmm_max(src, dst, n_out, msg_op, x=None, w=None):
    src, dst = _i64(src), _i64(dst)
    x2, w2, feat, ww = _prep(msg_op, src, x, w)
    out = np.empty((n_out, feat), dtype=np.float32)
    arg = np.empty((n_out, feat), dtype=np.int64)
    lib().oracle_spmm_max(_I(n_out), _I(len(src)), _p(src), _p(dst), ctypes.c_int(_MSG[msg_op]),
                          _p(x2), _p(w2), _I(feat), _I(ww), _p(out), _p(arg))
    return out, arg


def spmm_grad_x(src, dst, n_in, dy, w=None):
    src, dst = _i64(src), _i64(dst)
    dy2 = _f32(dy).reshape(len(dy), -1)
    w2 = None if w is None else _f32(w).reshape(len(src), -1)
    out = np.empty((n_in, dy2.shape[1]), dtype=np.float32)
    lib().oracle_spmm_grad_x(_I(n_in), _I(len(src)), _p(src), _p(dst), _p(dy2), _p(w2),
                             _I(dy2.shape[1]), _I(1 if w2 is None else w2.shape[1]), _p(out))
    return out


def degree_bucketing(msg_ids, vids, recv_ids, id_bound):
    msg_ids, vids, recv_ids = _i64(msg_ids), _i64(vids), _i64(recv_ids)
    m, r = len(vids), len(recv_ids)
    degs = np.empty(id_bound + 1, dtype=np.int64)
    nsec = np.empty(id_bound + 1, dtype=np.int64)
    msec = np.empty(id_bound + 1, dtype=np.int64)
    nids = np.empty(m + r, dtype=np.int64)
    mids = np.empty(max(m, 1), dtype=np.int64)
    counts = np.zeros(3, dtype=np.int64)
    lib().oracle_degree_bucketing(_I(m), _I(r), _I(id_bound), _p(msg_ids), _p(vids), _p(recv_ids),
                                  _p(degs), _p(nids), _p(nsec), _p(mids), _p(msec), _p(counts))
    b, k, bnz = counts.tolist()
    return degs[:b].copy(), nids[:k].copy(), nsec[:b].copy(), mids[:m].copy(), msec[:bnz].copy()


def edge_softmax(dst, score, n):
    dst = _i64(dst)
    s = _f32(score).reshape(len(dst), -1)
    out = np.empty_like(s)
    lib().oracle_edge_softmax(_I(n), _I(len(dst)), _p(dst), _I(s.shape[1]), _p(s), _p(out))
    return out.reshape(np.shape(score))
