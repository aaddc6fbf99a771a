"""TEST INFRASTRUCTURE (oracle): torch restatement of the non-rigid deformation forward, HashGridwithMLP.forward
(models/deformer/non_rigid.py:226-286): AABB.normalize (utils/dataset_utils.py:66-71), HashGrid (models/network_utils.py:
329-344), VanillaCondMLP (models/network_utils.py:182-249), delta application (non_rigid.py:256-282).

The hash-grid encoding is tiny-cuda-nn's (unvendored, unpinned dependency: README.md:45-46): the published multiresolution
hash encoding of Mueller et al. 2022 as implemented by tcnn's GridEncoding (scale = base * s^l - 1, res = ceil(scale) + 1,
pos = x * scale + 0.5, dense index while the level fits, else the xor-prime hash, index mod level size) is restated
here -- PARITY UNPINNED for this piece (no reference outputs available).  Only tests/ and __graft_entry__ import this."""
import math

import numpy as np
import torch

PRIMES = (1, 2654435761, 805459861)


def grid_levels(n_levels=16, base_resolution=16, per_level_scale=None, max_resolution=2048, log2_hashmap_size=16):
    """-> (level_offset [L+1], level_res [L], level_scale [L]) exactly as tcnn sizes the levels."""
    if per_level_scale is None:  # models/network_utils.py:332-336
        per_level_scale = float(np.exp(np.log(max_resolution / base_resolution) / (n_levels - 1)))
    off, res_l, scale_l = [0], [], []
    for l in range(n_levels):
        scale = np.float32(np.exp2(np.float32(l) * np.log2(np.float32(per_level_scale))) * np.float32(base_resolution) - np.float32(1.0))
        res = int(math.ceil(float(scale))) + 1
        n = min(res ** 3, 1 << 32)
        n = (n + 7) // 8 * 8
        n = min(n, 1 << log2_hashmap_size)
        off.append(off[-1] + n)
        res_l.append(res)
        scale_l.append(float(scale))
    return np.asarray(off, np.uint32), np.asarray(res_l, np.uint32), np.asarray(scale_l, np.float32)


def hash_encode(x01, table, level_offset, level_res, level_scale):
    """x01 [P,3] in [0,1]; table [n_entries, 2] -> [P, 2L]."""
    P = x01.shape[0]
    outs = []
    for l in range(len(level_res)):
        scale, res = float(level_scale[l]), int(level_res[l])
        off, size = int(level_offset[l]), int(level_offset[l + 1]) - int(level_offset[l])
        pos = x01 * scale + 0.5
        g = torch.floor(pos)
        w = pos - g
        g = g.to(torch.int64)
        acc = torch.zeros(P, table.shape[1], dtype=x01.dtype)
        for c in range(8):
            d = [(c >> k) & 1 for k in range(3)]
            wgt = torch.ones(P, dtype=x01.dtype)
            for k in range(3):
                wgt = wgt * (w[:, k] if d[k] else 1.0 - w[:, k])
            p = [(g[:, k] + d[k]) & 0xFFFFFFFF for k in range(3)]
            stride, dense_ok, idx = 1, True, torch.zeros(P, dtype=torch.int64)
            for k in range(3):
                if stride <= size:
                    idx = (idx + p[k] * stride) & 0xFFFFFFFF
                    stride *= res
            if size < stride:
                idx = (p[0] * PRIMES[0]) & 0xFFFFFFFF
                idx = idx ^ ((p[1] * PRIMES[1]) & 0xFFFFFFFF) ^ ((p[2] * PRIMES[2]) & 0xFFFFFFFF)
            idx = idx % size
            acc = acc + wgt[:, None] * table[off + idx]
        outs.append(acc)
    return torch.cat(outs, 1)


def _lrelu(h, mask=None):
    """LeakyReLU(0.01).  `mask` (optional, bool like h): the branch each unit takes, prescribed from outside instead of h > 0 --
    lets a gradient test use the SAME branches as the implementation under test for the few units whose pre-activation is
    within rounding of zero (their derivative jumps between 1 and 0.01)."""
    return torch.nn.functional.leaky_relu(h, 0.01) if mask is None else torch.where(mask, h, 0.01 * h)


def nonrigid_forward(xyz, scaling, rotation, aabb_min, aabb_max, table, levels, cond, W, b, features=None, masks=None):
    """W = [W0 [H, F+C], W1, W2, W3 [n_out, H]], b likewise.  Returns dict(xyz, scaling, rotation, deltas, features).
    `features` (optional) replaces the hash-grid encoding (to check the MLP on its own); `masks`: see _lrelu."""
    xn = 2.0 * ((xyz - aabb_min) / (aabb_max - aabb_min)) - 1.0  # AABB.normalize(sym=True)
    feat = hash_encode((xn + 1.0) * 0.5, table, *levels) if features is None else features
    h = torch.cat([feat, cond.reshape(1, -1).expand(xyz.shape[0], -1)], 1)  # cond_in: [0]
    for l in range(4):
        h = h @ W[l].T + b[l]
        if l < 3:
            h = _lrelu(h, None if masks is None else masks[l])
    d = h
    q1 = torch.cat([torch.ones_like(d[:, 6:7]), d[:, 7:10]], 1)  # q1[:, 0] = 1 (non_rigid.py:276-277)
    aw, ax, ay, az = q1.unbind(1)
    bw, bx, by, bz = rotation.unbind(1)
    rot = torch.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                       aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], 1)
    return dict(xyz=xyz + d[:, :3], scaling=scaling + d[:, 3:6], rotation=rot, deltas=d, features=feat)


def hierarchical_softmax(x):
    """models/deformer/rigid.py:85-129, statement by statement (x [P,25] -> [P,24])."""
    sig = torch.sigmoid(x)
    sm = lambda t: torch.softmax(t, dim=-1)
    p = torch.ones(x.shape[0], 24, dtype=x.dtype)
    p[:, [1, 2, 3]] = sig[:, [0]] * sm(x[:, [1, 2, 3]])
    p[:, [0]] = 1 - sig[:, [0]]
    p[:, [4, 5, 6]] = p[:, [1, 2, 3]] * sig[:, [4, 5, 6]]
    p[:, [1, 2, 3]] = p[:, [1, 2, 3]] * (1 - sig[:, [4, 5, 6]])
    p[:, [7, 8, 9]] = p[:, [4, 5, 6]] * sig[:, [7, 8, 9]]
    p[:, [4, 5, 6]] = p[:, [4, 5, 6]] * (1 - sig[:, [7, 8, 9]])
    p[:, [10, 11]] = p[:, [7, 8]] * sig[:, [10, 11]]
    p[:, [7, 8]] = p[:, [7, 8]] * (1 - sig[:, [10, 11]])
    p[:, [12, 13, 14]] = p[:, [9]] * sig[:, [24]] * sm(x[:, [12, 13, 14]])
    p[:, [9]] = p[:, [9]] * (1 - sig[:, [24]])
    p[:, [15]] = p[:, [12]] * sig[:, [15]]
    p[:, [12]] = p[:, [12]] * (1 - sig[:, [15]])
    p[:, [16, 17]] = p[:, [13, 14]] * sig[:, [16, 17]]
    p[:, [13, 14]] = p[:, [13, 14]] * (1 - sig[:, [16, 17]])
    p[:, [18, 19]] = p[:, [16, 17]] * sig[:, [18, 19]]
    p[:, [16, 17]] = p[:, [16, 17]] * (1 - sig[:, [18, 19]])
    p[:, [20, 21]] = p[:, [18, 19]] * sig[:, [20, 21]]
    p[:, [18, 19]] = p[:, [18, 19]] * (1 - sig[:, [20, 21]])
    p[:, [22, 23]] = p[:, [20, 21]] * sig[:, [22, 23]]
    p[:, [20, 21]] = p[:, [20, 21]] * (1 - sig[:, [22, 23]])
    return p


def skinning_weights_forward(xyz, aabb_min, aabb_max, W, b, masks=None):
    """SkinningField.get_forward_transform's weights (rigid.py:160-171, 189-196, non-distilled): AABB.normalize(sym) ->
    VanillaCondMLP(3 -> H x4 -> n_out), LeakyReLU(0.01) -> hierarchical softmax (25 logits) / softmax (24)."""
    h = 2.0 * ((xyz - aabb_min) / (aabb_max - aabb_min)) - 1.0
    for l in range(len(W)):
        h = h @ W[l].T + b[l]
        if l < len(W) - 1:
            h = _lrelu(h, None if masks is None else masks[l])
    w = hierarchical_softmax(h) if h.shape[1] == 25 else torch.softmax(h, -1)
    return w, h
