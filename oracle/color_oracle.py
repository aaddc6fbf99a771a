"""TEST INFRASTRUCTURE (oracle): torch-CPU restatement of the colour MLP texture, ColorMLP.compose_input / forward
(models/texture/texture.py:74-125) with VanillaCondMLP (models/network_utils.py:182-249) and `eval_sh_bases`
(utils/sh_utils.py:115-161).  The reference class itself cannot be imported here (models/network_utils.py:4 imports
tinycudann), so the composition is restated statement by statement; `eval_sh_bases` IS importable on CPU and
tests/test_oracle_color.py pins `sh_bases` against it (golden vector committed under tests/golden/).  Runs in fp32
(parity) or fp64 (truth for the tensor-core path's tolerance).  Only tests/ and __graft_entry__ import this."""
import torch
import torch.nn.functional as F

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]


def sh_bases(deg, dirs):
    """eval_sh_bases(deg, dirs), utils/sh_utils.py:115-149 (deg <= 3): [..., 3] unit directions -> [..., (deg+1)^2]."""
    assert 0 <= deg <= 3
    x, y, z = dirs.unbind(-1)
    out = [torch.full_like(x, C0)]
    if deg > 0:
        out += [-C1 * y, C1 * z, -C1 * x]
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            out += [C2[0] * xy, C2[1] * yz, C2[2] * (2.0 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy)]
            if deg > 2:
                out += [C3[0] * y * (3 * xx - yy), C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy),
                        C3[3] * z * (2 * zz - 3 * xx - 3 * yy), C3[4] * x * (4 * zz - xx - yy), C3[5] * z * (xx - yy),
                        C3[6] * x * (xx - 3 * yy)]
    return torch.stack(out, -1)


def compose_input(features, xyz, T_fwd, camera_center, non_rigid_feature, latent, sh_degree=3, cano_view_dir=True,
                  view_noise=None):
    """texture.py:74-118 (use_xyz / use_cov / use_normal off, as in configs/texture/shallow_mlp.yaml)."""
    n = features.shape[0]
    parts = [features]
    if sh_degree > 0:
        dir_pp = xyz - camera_center.reshape(1, 3).repeat(n, 1)
        if cano_view_dir:
            R_bwd = T_fwd.reshape(n, 4, 4)[:, :3, :3].transpose(1, 2)
            dir_pp = torch.matmul(R_bwd, dir_pp.unsqueeze(-1)).squeeze(-1)
            if view_noise is not None:
                dir_pp = torch.matmul(dir_pp, view_noise)
        dir_n = dir_pp / (dir_pp.norm(dim=1, keepdim=True) + 1e-12)
        parts.append(sh_bases(sh_degree, dir_n)[..., 1:])
    if non_rigid_feature is not None:
        parts.append(non_rigid_feature)
    if latent is not None:
        parts.append(latent.reshape(1, -1).expand(n, -1))
    return torch.cat(parts, 1)


def color_forward(features, xyz, T_fwd, camera_center, non_rigid_feature, latent, Ws, bs, sh_degree=3,
                  cano_view_dir=True, view_noise=None, masks=None):
    """-> colors [P,3]: sigmoid(lin2(lrelu(lin1(lrelu(lin0(input)))))).  `masks` (optional): the LeakyReLU branch of every
    hidden unit prescribed from outside (see oracle/nonrigid_oracle.py:_lrelu)."""
    h = compose_input(features, xyz, T_fwd, camera_center, non_rigid_feature, latent, sh_degree, cano_view_dir, view_noise)
    for l in range(len(Ws)):
        h = F.linear(h, Ws[l], bs[l])
        if l < len(Ws) - 1:
            h = F.leaky_relu(h, 0.01) if masks is None else torch.where(masks[l], h, 0.01 * h)
    return torch.sigmoid(h)


def synthetic_case(P, seed=0, feature_dim=32, sh_degree=3, non_rigid_dim=16, latent_dim=16, hidden=64, dtype=torch.float32):
    """Seeded inputs + nn.Linear-style initialised weights of the shallow_mlp configuration."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g, dtype=torch.float64)
    d_in = feature_dim + (sh_degree + 1) ** 2 - 1 + non_rigid_dim + latent_dim
    dims = [d_in, hidden, hidden, 3]
    Ws, bs = [], []
    for l in range(3):
        bound = 1.0 / dims[l] ** 0.5
        Ws.append(((torch.rand(dims[l + 1], dims[l], generator=g, dtype=torch.float64) * 2 - 1) * bound * 3.0).to(dtype))
        bs.append(((torch.rand(dims[l + 1], generator=g, dtype=torch.float64) * 2 - 1) * bound).to(dtype))
    # per-Gaussian rigid transforms: random rotations (QR) + translations
    q, _ = torch.linalg.qr(r(P, 3, 3))
    T = torch.eye(4, dtype=torch.float64).repeat(P, 1, 1)
    T[:, :3, :3] = q
    T[:, :3, 3] = 0.3 * r(P, 3)
    ang = torch.tensor(0.4, dtype=torch.float64)
    noise = torch.tensor([[torch.cos(ang), -torch.sin(ang), 0.0], [torch.sin(ang), torch.cos(ang), 0.0], [0.0, 0.0, 1.0]], dtype=torch.float64)
    return dict(features=r(P, feature_dim).to(dtype), xyz=(0.5 * r(P, 3)).to(dtype), T_fwd=T.reshape(P, 16).to(dtype),
                camera_center=torch.tensor([0.2, 0.9, 3.5], dtype=dtype), non_rigid_feature=(0.5 * r(P, non_rigid_dim)).to(dtype) if non_rigid_dim else None,
                latent=(0.5 * r(latent_dim)).to(dtype) if latent_dim else None, Ws=Ws, bs=bs, view_noise=noise.to(dtype),
                sh_degree=sh_degree)
