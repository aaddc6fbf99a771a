"""Drop-ins for the reference's three model registries, same function names, same `cfg.name` keys, same (cfg, metadata)
signature: get_rigid_deform (models/deformer/rigid.py:244-251), get_non_rigid_deform (models/deformer/non_rigid.py:302-310),
get_texture (models/texture/texture.py:128-134).  `cfg` is the reference's OmegaConf node (attribute access and .get());
`metadata` the dataset metadata dict (smpl_verts, skinning_weights, aabb, frame_dict).  The returned modules are the fused
B200 stages consumed by FusedGaussianConverter(rigid, non_rigid, pose_encoder, texture=...):

    rigid     = get_rigid_deform(cfg.model.deformer.rigid, metadata)          # smpl_nn | skinning_field
    non_rigid = get_non_rigid_deform(cfg.model.deformer.non_rigid, metadata)  # identity (None) | hashgrid
    texture   = get_texture(cfg.model.texture, metadata)                      # sh2rgb (None: SH inside fused_pose) | mlp
    converter = FusedGaussianConverter(rigid, non_rigid, getattr(non_rigid, "pose_encoder", None), texture=texture)

Keys outside the hot path named by BASELINE.json's north_star (rigid `identity`, non-rigid `mlp` / `hannw_mlp`: ablation models
of the reference) raise NotImplementedError with the key, so a configuration that needs them fails loudly."""
import numpy as np
import torch


def _get(cfg, key, default=None):
    if hasattr(cfg, "get"):
        try:
            return cfg.get(key, default)
        except TypeError:
            pass
    return getattr(cfg, key, default)


def _aabb_bounds(aabb):
    """(min [3], max [3]) from the reference's AABB object (utils/dataset_utils.py:45-76: coord_min / coord_max) or a pair."""
    if hasattr(aabb, "coord_min"):
        lo, hi = aabb.coord_min, aabb.coord_max
    else:
        lo, hi = aabb
    f = lambda v: [float(x) for x in np.asarray(v.detach().cpu() if torch.is_tensor(v) else v, dtype=np.float64).reshape(-1)[:3]]
    return f(lo), f(hi)


def get_rigid_deform(cfg, metadata):
    from .modules import SMPLNNFused
    from .skinning import SkinningFieldFused
    name = cfg.name
    if name == "smpl_nn":  # rigid.py:35-71
        return SMPLNNFused(metadata["smpl_verts"], metadata["skinning_weights"])
    if name == "skinning_field":  # rigid.py:131-236, configs/rigid/skinning_field.yaml
        net = _get(cfg, "skinning_network")
        hidden = int(_get(net, "n_neurons", 128)) if net is not None else 128
        layers = int(_get(net, "n_hidden_layers", 4)) if net is not None else 4
        if hidden != 128 or layers != 4 or _get(cfg, "distill", False):
            raise NotImplementedError("skinning_field: the fused stage covers the reference's configuration (128 x 4, distill: false)")
        lo, hi = _aabb_bounds(metadata["aabb"])
        return SkinningFieldFused(d_out=int(_get(cfg, "d_out", 25)), hidden=hidden, aabb_min=lo, aabb_max=hi)
    raise NotImplementedError(f"rigid deformer '{name}' is outside the fused hot path (keys: smpl_nn, skinning_field)")


def get_non_rigid_deform(cfg, metadata):
    from .nonrigid import HashGridMLPFused
    from .pose_encoder import HierarchicalPoseEncoderFused
    name = cfg.name
    if name == "identity":  # non_rigid.py:22-29: FusedGaussianConverter(non_rigid=None)
        return None
    if name == "hashgrid":  # non_rigid.py:203-300, configs/non_rigid/hashgrid.yaml
        if int(_get(cfg, "latent_dim", 0)) != 0:
            raise NotImplementedError("hashgrid: latent_dim > 0 is not part of the fused stage (reference default: 0)")
        hg, mlp = cfg.hashgrid, cfg.mlp
        if int(_get(hg, "n_features_per_level", 2)) != 2 or int(_get(mlp, "n_neurons", 128)) != 128 or int(_get(mlp, "n_hidden_layers", 3)) != 3:
            raise NotImplementedError("hashgrid: the fused stage covers 2 features per level and a 128 x 3 MLP (the reference's configuration)")
        pe = _get(cfg, "pose_encoder")
        encoder = HierarchicalPoseEncoderFused(rel_joints=bool(_get(pe, "rel_joints", False)), dim_per_joint=int(_get(pe, "dim_per_joint", 6)))
        lo, hi = _aabb_bounds(metadata["aabb"])
        m = HashGridMLPFused(n_cond=int(encoder.n_output_dims), feature_dim=int(_get(cfg, "feature_dim", 0)), aabb_min=lo, aabb_max=hi,
                             n_levels=int(_get(hg, "n_levels", 16)), base_resolution=int(_get(hg, "base_resolution", 16)),
                             max_resolution=int(_get(hg, "max_resolution", 2048)), log2_hashmap_size=int(_get(hg, "log2_hashmap_size", 16)),
                             hidden=128, delay=int(_get(cfg, "delay", 0)))
        m.pose_encoder = encoder  # HashGridwithMLP owns its HierarchicalPoseEncoder (non_rigid.py:206)
        return m
    raise NotImplementedError(f"non-rigid deformer '{name}' is outside the fused hot path (keys: identity, hashgrid)")


def get_texture(cfg, metadata):
    from .texture import ColorMLPFused
    name = cfg.name
    if name == "sh2rgb":  # texture.py:17-38: evaluated inside the fused posing kernel (FusedGaussianConverter(texture=None))
        return None
    if name == "mlp":  # texture.py:40-125, configs/texture/shallow_mlp.yaml
        if _get(cfg, "use_xyz", False) or _get(cfg, "use_cov", False) or _get(cfg, "use_normal", False):
            raise NotImplementedError("texture mlp: use_xyz / use_cov / use_normal are off in the reference's configurations and not fused")
        mlp = cfg.mlp
        if int(_get(mlp, "n_neurons", 64)) != 64 or int(_get(mlp, "n_hidden_layers", 2)) != 2:
            raise NotImplementedError("texture mlp: the fused stage covers the 64 x 2 shallow MLP (configs/texture/shallow_mlp.yaml)")
        latent_dim = int(_get(cfg, "latent_dim", 0))
        n_frames = len(metadata["frame_dict"]) if latent_dim > 0 else 1
        return ColorMLPFused(feature_dim=int(cfg.feature_dim), sh_degree=int(_get(cfg, "sh_degree", 0)), non_rigid_dim=int(_get(cfg, "non_rigid_dim", 0)),
                             latent_dim=latent_dim, n_frames=n_frames, hidden=64, cano_view_dir=bool(_get(cfg, "cano_view_dir", False)))
    raise NotImplementedError(f"texture '{name}' is not known (keys: sh2rgb, mlp)")
