"""The registry drop-ins (avatar_pose.registry) take the reference's own configuration nodes: the YAML files under
/root/reference/configs are parsed where present (skipped otherwise) and the returned fused modules carry the configured
shapes.  Reference: rigid.py:244-251, non_rigid.py:302-310, texture.py:128-134."""
import os

import numpy as np
import pytest
import yaml

CFG = "/root/reference/configs"


class Node(dict):
    """Minimal OmegaConf stand-in: attribute access + .get()."""
    __getattr__ = dict.__getitem__

    @staticmethod
    def wrap(v):
        return Node({k: Node.wrap(x) for k, x in v.items()}) if isinstance(v, dict) else v


def load(rel):
    with open(os.path.join(CFG, rel)) as f:
        return Node.wrap(yaml.safe_load(f))


@pytest.mark.skipif(not os.path.isdir(CFG), reason="reference configs not present")
def test_registry_builds_default_configuration():
    from avatar_pose.registry import get_rigid_deform, get_non_rigid_deform, get_texture
    from avatar_pose.modules import SMPLNNFused, FusedGaussianConverter
    from avatar_pose.skinning import SkinningFieldFused
    from avatar_pose.nonrigid import HashGridMLPFused
    from avatar_pose.texture import ColorMLPFused
    rng = np.random.default_rng(0)
    aabb = type("AABB", (), {"coord_min": np.array([-1.0, -1.2, -0.5]), "coord_max": np.array([1.0, 1.0, 0.6])})()
    metadata = {"smpl_verts": rng.normal(size=(6890, 3)).astype(np.float32), "skinning_weights": rng.random((6890, 24)).astype(np.float32),
                "aabb": aabb, "frame_dict": {i: i for i in range(12)}}
    rigid_cfg = load("rigid/skinning_field.yaml").model.deformer.rigid
    nr_cfg = load("non_rigid/hashgrid.yaml").model.deformer.non_rigid
    nr_cfg["feature_dim"] = 16  # ${model.texture.non_rigid_dim}
    tex_cfg = load("texture/shallow_mlp.yaml").model.texture
    tex_cfg["feature_dim"] = 32  # ${model.gaussian.feature_dim}
    rigid = get_rigid_deform(rigid_cfg, metadata)
    assert isinstance(rigid, SkinningFieldFused) and rigid.d_out == 25 and rigid.lin4.weight.shape == (25, 128)
    assert np.allclose(rigid.aabb_min.numpy(), aabb.coord_min)
    nr = get_non_rigid_deform(nr_cfg, metadata)
    assert isinstance(nr, HashGridMLPFused) and nr.delay == 3000 and nr.n_out == 26 and nr.lin0.weight.shape == (128, 32 + 144)
    assert nr.pose_encoder.n_output_dims == 144
    tex = get_texture(tex_cfg, metadata)
    assert isinstance(tex, ColorMLPFused) and tex.lin0.weight.shape == (64, 32 + 15 + 16 + 16) and tex.latent.num_embeddings == 12
    assert isinstance(get_rigid_deform(load("rigid/smpl_nn.yaml").model.deformer.rigid, metadata), SMPLNNFused)
    assert get_non_rigid_deform(load("non_rigid/identity.yaml").model.deformer.non_rigid, metadata) is None
    assert get_texture(load("texture/sh.yaml").model.texture, metadata) is None
    conv = FusedGaussianConverter(rigid, nr, nr.pose_encoder, texture=tex)
    assert len(list(conv.parameters())) > 20
    with pytest.raises(NotImplementedError):
        get_non_rigid_deform(load("non_rigid/mlp.yaml").model.deformer.non_rigid, metadata)
    with pytest.raises(NotImplementedError):
        get_rigid_deform(load("rigid/identity.yaml").model.deformer.rigid, metadata)
