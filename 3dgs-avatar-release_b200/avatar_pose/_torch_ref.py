"""Differentiable PyTorch (GPU) re-evaluation of the two MLP stages, used ONLY by the backward of
`avatar_pose.nonrigid` / `avatar_pose.skinning` in this round: the forward runs on the hand-written tcgen05 kernels; the
backward re-evaluates the same math with torch ops under autograd (cuBLAS GEMMs, index_add for the hash table) to produce
parameter and input gradients.  A native backward (dW accumulated in TMEM across the tiles of a persistent CTA) is the
next step.  This is GPU code, not a CPU fallback; it is never used for forward results."""
import torch
import torch.nn.functional as F

PRIMES = (1, 2654435761, 805459861)


def hash_encode(x01, table, level_offset, level_res, level_scale):
    """Multiresolution hash encoding (tiny-cuda-nn GridEncoding semantics, see gs_nonrigid.cu): x01 [P,3] -> [P,2L]."""
    P = x01.shape[0]
    outs = []
    for l in range(len(level_res)):
        scale, res = float(level_scale[l]), int(level_res[l])
        off, size = int(level_offset[l]), int(level_offset[l + 1]) - int(level_offset[l])
        pos = x01 * scale + 0.5
        g = torch.floor(pos)
        w = pos - g
        g = g.detach().to(torch.int64)
        acc = torch.zeros(P, table.shape[1], dtype=x01.dtype, device=x01.device)
        for c in range(8):
            d = [(c >> k) & 1 for k in range(3)]
            wgt = torch.ones(P, dtype=x01.dtype, device=x01.device)
            for k in range(3):
                wgt = wgt * (w[:, k] if d[k] else 1.0 - w[:, k])
            p = [(g[:, k] + d[k]) & 0xFFFFFFFF for k in range(3)]
            stride, idx = 1, torch.zeros(P, dtype=torch.int64, device=x01.device)
            for k in range(3):
                if stride <= size:
                    idx = (idx + p[k] * stride) & 0xFFFFFFFF
                    stride *= res
            if size < stride:
                idx = ((p[0] * PRIMES[0]) & 0xFFFFFFFF) ^ ((p[1] * PRIMES[1]) & 0xFFFFFFFF) ^ ((p[2] * PRIMES[2]) & 0xFFFFFFFF)
            idx = idx % size
            acc = acc + wgt[:, None] * table[off + idx]
        outs.append(acc)
    return torch.cat(outs, 1)


def mlp(h, Ws, bs):
    for l in range(len(Ws)):
        h = F.linear(h, Ws[l], bs[l])
        if l < len(Ws) - 1:
            h = F.leaky_relu(h, 0.01)
    return h


def nonrigid(xyz, scaling, rotation, cond, table, Ws, bs, aabb_min, aabb_max, levels):
    """-> (xyz', scaling', rotation', deltas): non_rigid.py:248-282 (scale_offset logit, rot_offset mult)."""
    xn = 2.0 * ((xyz - aabb_min) / (aabb_max - aabb_min)) - 1.0
    feat = hash_encode((xn + 1.0) * 0.5, table, *levels)
    d = mlp(torch.cat([feat, cond.reshape(1, -1).expand(xyz.shape[0], -1)], 1), Ws, bs)
    q1 = torch.cat([torch.ones_like(d[:, 6:7]), d[:, 7:10]], 1)
    aw, ax, ay, az = q1.unbind(1)
    bw, bx, by, bz = rotation.unbind(1)
    rot = torch.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                       aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], 1)
    return xyz + d[:, :3], scaling + d[:, 3:6], rot, d


def hierarchical_softmax(x):
    """rigid.py:85-129 without in-place writes (autograd friendly): x [P,25] -> [P,24]."""
    s = torch.sigmoid(x)
    sm = lambda t: torch.softmax(t, dim=-1)
    p = [None] * 24
    a = s[:, 0:1] * sm(x[:, 1:4])
    p[0] = 1 - s[:, 0]
    l1 = [a[:, k] for k in range(3)]                                   # joints 1,2,3
    l2 = [l1[k] * s[:, 4 + k] for k in range(3)]                       # 4,5,6
    for k in range(3):
        p[1 + k] = l1[k] * (1 - s[:, 4 + k])
    l3 = [l2[k] * s[:, 7 + k] for k in range(3)]                       # 7,8,9
    for k in range(3):
        p[4 + k] = l2[k] * (1 - s[:, 7 + k])
    l4 = [l3[k] * s[:, 10 + k] for k in range(2)]                      # 10,11
    for k in range(2):
        p[7 + k] = l3[k] * (1 - s[:, 10 + k])
        p[10 + k] = l4[k]
    b = (l3[2] * s[:, 24])[:, None] * sm(x[:, 12:15])                  # 12,13,14 from joint 9 gated by logit 24
    p[9] = l3[2] * (1 - s[:, 24])
    n12, n13, n14 = b[:, 0], b[:, 1], b[:, 2]
    p[15] = n12 * s[:, 15]
    p[12] = n12 * (1 - s[:, 15])
    c16 = [n13 * s[:, 16], n14 * s[:, 17]]
    p[13], p[14] = n13 * (1 - s[:, 16]), n14 * (1 - s[:, 17])
    c18 = [c16[k] * s[:, 18 + k] for k in range(2)]
    p[16], p[17] = c16[0] * (1 - s[:, 18]), c16[1] * (1 - s[:, 19])
    c20 = [c18[k] * s[:, 20 + k] for k in range(2)]
    p[18], p[19] = c18[0] * (1 - s[:, 20]), c18[1] * (1 - s[:, 21])
    c22 = [c20[k] * s[:, 22 + k] for k in range(2)]
    p[20], p[21] = c20[0] * (1 - s[:, 22]), c20[1] * (1 - s[:, 23])
    p[22], p[23] = c22[0], c22[1]
    return torch.stack(p, 1)


def skinning_weights(xyz, Ws, bs, aabb_min, aabb_max):
    h = mlp(2.0 * ((xyz - aabb_min) / (aabb_max - aabb_min)) - 1.0, Ws, bs)
    return hierarchical_softmax(h) if h.shape[1] == 25 else torch.softmax(h, -1)


_SH_C1 = 0.4886025119029199
_SH_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
_SH_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
          1.445305721320277, -0.5900435899266435)


def sh_bases_1(deg, d):
    """eval_sh_bases(deg, d)[..., 1:] (utils/sh_utils.py:115-149), autograd friendly."""
    x, y, z = d.unbind(-1)
    out = []
    if deg > 0:
        out += [-_SH_C1 * y, _SH_C1 * z, -_SH_C1 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        out += [_SH_C2[0] * xy, _SH_C2[1] * yz, _SH_C2[2] * (2.0 * zz - xx - yy), _SH_C2[3] * xz, _SH_C2[4] * (xx - yy)]
    if deg > 2:
        out += [_SH_C3[0] * y * (3 * xx - yy), _SH_C3[1] * xy * z, _SH_C3[2] * y * (4 * zz - xx - yy),
                _SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), _SH_C3[4] * x * (4 * zz - xx - yy), _SH_C3[5] * z * (xx - yy),
                _SH_C3[6] * x * (xx - 3 * yy)]
    return torch.stack(out, -1)


def color_mlp(features, xyz, T_fwd, camera_center, non_rigid_feature, latent, Ws, bs, sh_degree, cano_view_dir, view_noise):
    """ColorMLP.forward (texture.py:74-125) for the backward re-evaluation of avatar_pose.texture."""
    n = features.shape[0]
    parts = [features]
    if sh_degree > 0:
        d = xyz - camera_center.reshape(1, 3)
        if cano_view_dir:
            d = torch.matmul(T_fwd.reshape(n, 4, 4)[:, :3, :3].transpose(1, 2), d.unsqueeze(-1)).squeeze(-1)
            if view_noise is not None:
                d = torch.matmul(d, view_noise)
        parts.append(sh_bases_1(sh_degree, d / (d.norm(dim=1, keepdim=True) + 1e-12)))
    if non_rigid_feature is not None:
        parts.append(non_rigid_feature)
    if latent is not None:
        parts.append(latent.reshape(1, -1).expand(n, -1))
    return torch.sigmoid(mlp(torch.cat(parts, 1), Ws, bs))
