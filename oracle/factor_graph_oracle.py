"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's factor-graph staging, the checker for
monst3r-slam_b200/factor_graph.py and csrc/edge_store.cu.  Nothing in the product imports this file.

Follows mast3r_slam/global_opt.py:54-110 (the part of FactorGraph.add_factors after the ViT matcher,
get_unique_kf_idx, prep_two_way_edges) and mast3r_slam/geometry.py:37-42,107-123
(constrain_points_to_ray = get_pixel_coords + backproject).

Parity pinned: tests/test_factor_graph_oracle.py checks every function against
tests/golden/factor_graph.npz, which tests/golden/make_golden_factor_graph.py produced by running the
reference's own python (imported from /root/reference, only the ViT matcher stubbed).
"""
import numpy as np


class GraphState:
    """the eight growing tensors of FactorGraph.__init__ (global_opt.py:17-25)"""

    def __init__(self, n):
        self.n = n
        self.ii = np.zeros((0,), np.int64)
        self.jj = np.zeros((0,), np.int64)
        self.idx_ii2jj = np.zeros((0, n), np.int64)
        self.idx_jj2ii = np.zeros((0, n), np.int64)
        self.valid_match_j = np.zeros((0, n, 1), bool)
        self.valid_match_i = np.zeros((0, n, 1), bool)
        self.Q_ii2jj = np.zeros((0, n, 1), np.float32)
        self.Q_jj2ii = np.zeros((0, n, 1), np.float32)


def edge_q(Q_src, idx, Q_dst):
    """global_opt.py:54-58: Qj = sqrt(Qii[batch_inds, idx_i2j] * Qji), float32 throughout"""
    b = np.arange(idx.shape[0])[:, None]
    return np.sqrt(Q_src.astype(np.float32)[b, idx] * Q_dst.astype(np.float32)).astype(np.float32)


def add_factors_tail(st, ii, jj, matches, Q_conf, min_match_frac, is_reloc=False):
    """global_opt.py:54-99 -> the method's return value; `st` is updated like the reference's attributes"""
    idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij = matches
    Qj = edge_q(Qii, idx_i2j, Qji)  # :57
    Qi = edge_q(Qjj, idx_j2i, Qij)  # :58
    valid_j = valid_match_j & (Qj > np.float32(Q_conf))  # :60-63
    valid_i = valid_match_i & (Qi > np.float32(Q_conf))
    nj = valid_j.shape[1] * valid_j.shape[2]
    ni = valid_i.shape[1] * valid_i.shape[2]
    # :66-67: int64 sum / int -> float32 true division in torch
    match_frac_j = valid_j.sum(axis=(1, 2)).astype(np.float32) / np.float32(nj)
    match_frac_i = valid_i.sum(axis=(1, 2)).astype(np.float32) / np.float32(ni)
    ii_t = np.asarray(ii, np.int64)
    jj_t = np.asarray(jj, np.int64)
    invalid = np.minimum(match_frac_j, match_frac_i) < np.float32(min_match_frac)  # :73
    consecutive = ii_t == (jj_t - 1)  # :74
    invalid = (~consecutive) & invalid  # :75
    if invalid.any() and is_reloc:  # :77-78
        return False
    keep = ~invalid
    st.ii = np.concatenate([st.ii, ii_t[keep]])
    st.jj = np.concatenate([st.jj, jj_t[keep]])
    st.idx_ii2jj = np.concatenate([st.idx_ii2jj, idx_i2j[keep]])
    st.idx_jj2ii = np.concatenate([st.idx_jj2ii, idx_j2i[keep]])
    st.valid_match_j = np.concatenate([st.valid_match_j, valid_match_j[keep]])
    st.valid_match_i = np.concatenate([st.valid_match_i, valid_match_i[keep]])
    st.Q_ii2jj = np.concatenate([st.Q_ii2jj, Qj[keep]])
    st.Q_jj2ii = np.concatenate([st.Q_jj2ii, Qi[keep]])
    return bool(keep.sum() > 0)  # :98-99


def get_unique_kf_idx(st):
    """:101-102"""
    return np.unique(np.concatenate([st.ii, st.jj]))


def prep_two_way_edges(st):
    """:104-110: directed edges = [all i->j, then all j->i]"""
    return (np.concatenate([st.ii, st.jj]), np.concatenate([st.jj, st.ii]),
            np.concatenate([st.idx_ii2jj, st.idx_jj2ii]), np.concatenate([st.valid_match_j, st.valid_match_i]),
            np.concatenate([st.Q_ii2jj, st.Q_jj2ii]))


def interleave_two_way(two_way):
    """reference order [fwd..., bwd...] -> the store's pair-interleaved order [fwd0, bwd0, fwd1, bwd1, ...]"""
    out = []
    for t in two_way:
        E = t.shape[0] // 2
        z = np.empty_like(t)
        z[0::2] = t[:E]
        z[1::2] = t[E:]
        out.append(z)
    return tuple(out)


def constrain_points_to_ray(img_size, Xs, K):
    """geometry.py:37-42 with get_pixel_coords (:117-123) and backproject (:107-114), float32"""
    h, w = img_size
    Xs = np.asarray(Xs, np.float32)
    K = np.asarray(K, np.float32)
    k = np.arange(h * w)
    u = (k % w).astype(np.float32)
    v = (k // w).astype(np.float32)
    shape = Xs.shape[:-1]
    u = np.broadcast_to(u, (Xs.size // (3 * h * w), h * w)).reshape(shape)
    v = np.broadcast_to(v, (Xs.size // (3 * h * w), h * w)).reshape(shape)
    z = Xs[..., 2]
    tmp1 = (u - K[0, 2]) / K[0, 0]
    tmp2 = (v - K[1, 2]) / K[1, 1]
    return np.stack([z * tmp1, z * tmp2, z * np.float32(1.0)], axis=-1).astype(np.float32)
