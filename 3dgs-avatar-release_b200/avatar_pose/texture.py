"""Colour MLP texture (forward on the tensor cores): host side of `color_forward` (include/gsplat_b200.h).  Mirrors
ColorMLP (models/texture/texture.py:40-125, registry key `mlp` of get_texture :128-134; configs/texture/shallow_mlp.yaml):
`forward(gaussians, camera)` -> colors_precomp [P,3].  `colors` is the tcgen05 forward; `colors_autograd` adds the
hand-written backward `color_backward` (tcgen05 dX / dW GEMMs over the activations the training forward kept + the adjoint of
compose_input): no torch ops, no cuBLAS.  No CPU fallback."""
import ctypes

import torch
import torch.nn as nn

from diff_gaussian_rasterization import _C as _gs

vp = ctypes.c_void_p


class ColorArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("feature_dim", ctypes.c_int32), ("sh_degree", ctypes.c_int32),
                ("non_rigid_dim", ctypes.c_int32), ("latent_dim", ctypes.c_int32), ("hidden", ctypes.c_int32),
                ("cano_view_dir", ctypes.c_int32), ("pad_", ctypes.c_int32),
                ("features", vp), ("xyz", vp), ("T_fwd", vp), ("camera_center", vp), ("view_noise", vp),
                ("non_rigid_feature", vp), ("latent", vp), ("W0", vp), ("b0", vp), ("W1", vp), ("b1", vp), ("W2", vp),
                ("b2", vp), ("colors", vp), ("inputs", vp)]


class ColorBackwardArgs(ctypes.Structure):
    _fields_ = [("fwd", ColorArgs), ("dL_dcolors", vp), ("dL_dfeatures", vp), ("dL_dxyz", vp), ("dL_dT_fwd", vp),
                ("dL_dnon_rigid_feature", vp), ("dL_dlatent", vp), ("dL_dW0", vp), ("dL_db0", vp), ("dL_dW1", vp), ("dL_db1", vp),
                ("dL_dW2", vp), ("dL_db2", vp)]


_configured = False


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.color_workspace_bytes.restype = ctypes.c_size_t
        L.color_workspace_bytes.argtypes = [ctypes.POINTER(ColorArgs)]
        L.color_forward.restype = ctypes.c_int
        L.color_forward.argtypes = [ctypes.POINTER(ColorArgs), vp, ctypes.c_size_t, vp]
        L.color_saved_bytes.restype = ctypes.c_size_t
        L.color_saved_bytes.argtypes = [ctypes.c_int32]
        L.color_forward_train.restype = ctypes.c_int
        L.color_forward_train.argtypes = [ctypes.POINTER(ColorArgs), vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]
        L.color_backward.restype = ctypes.c_int
        L.color_backward.argtypes = [ctypes.POINTER(ColorBackwardArgs), vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]
        _configured = True
    return L


class ColorMLPFused(nn.Module):
    def __init__(self, feature_dim=32, sh_degree=3, non_rigid_dim=16, latent_dim=16, n_frames=1, hidden=64,
                 cano_view_dir=True):
        super().__init__()
        self.feature_dim, self.sh_degree, self.non_rigid_dim, self.latent_dim = feature_dim, sh_degree, non_rigid_dim, latent_dim
        self.hidden, self.cano_view_dir = hidden, cano_view_dir
        d_in = feature_dim + ((sh_degree + 1) ** 2 - 1 if sh_degree > 0 else 0) + non_rigid_dim + latent_dim
        dims = [d_in, hidden, hidden, 3]  # n_hidden_layers: 2 (configs/texture/shallow_mlp.yaml)
        for l in range(3):
            setattr(self, f"lin{l}", nn.Linear(dims[l], dims[l + 1]))
        if latent_dim > 0:
            self.latent = nn.Embedding(n_frames, latent_dim)  # texture.py:66-67
        self._ws = None
        self._idx_cache = {}

    def params(self):
        return [t for l in range(3) for t in (getattr(self, f"lin{l}").weight, getattr(self, f"lin{l}").bias)]

    def latent_code(self, latent_idx):
        if self.latent_dim == 0:
            return None
        i, dev = min(int(latent_idx), self.latent.num_embeddings - 1), self.latent.weight.device
        idx = self._idx_cache.get((i, dev))  # device-resident index: a host -> device copy here could not be graph-captured
        if idx is None:
            idx = self._idx_cache[(i, dev)] = torch.full((1,), i, dtype=torch.long, device=dev)
        return self.latent(idx).reshape(-1)  # texture.py:107-114

    def _fill(self, a, features, xyz, T_fwd, camera_center, non_rigid_feature, latent, view_noise, keep):
        c = lambda t: None if t is None else t.detach().contiguous().float()
        P = features.shape[0]
        features, xyz, camera_center, non_rigid_feature, latent, view_noise = map(
            c, (features, xyz, camera_center, non_rigid_feature, latent, view_noise))
        T_fwd = None if T_fwd is None else c(T_fwd.reshape(P, 16))
        a.P, a.feature_dim, a.sh_degree, a.non_rigid_dim = P, self.feature_dim, self.sh_degree, self.non_rigid_dim
        a.latent_dim, a.hidden, a.cano_view_dir = self.latent_dim, self.hidden, int(self.cano_view_dir)
        p = _gs._ptr
        ws = [c(t) for t in self.params()]
        keep += ws + [features, xyz, T_fwd, camera_center, non_rigid_feature, latent, view_noise]
        a.W0, a.b0, a.W1, a.b1, a.W2, a.b2 = (p(t, "W") for t in ws)
        a.features, a.xyz, a.T_fwd = p(features, "features"), p(xyz, "xyz"), p(T_fwd, "T_fwd")
        a.camera_center, a.view_noise = p(camera_center, "camera_center"), p(view_noise, "view_noise")
        a.non_rigid_feature, a.latent = p(non_rigid_feature, "non_rigid_feature"), p(latent, "latent")

    def _workspace(self, L, a, dev):
        need = L.color_workspace_bytes(ctypes.byref(a))
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        return self._ws

    @torch.no_grad()
    def colors(self, features, xyz, T_fwd, camera_center, non_rigid_feature=None, latent=None, view_noise=None,
               return_inputs=False, _saved=None):
        """features [P,F], xyz [P,3] deformed means, T_fwd [P,4,4] or [P,16], camera_center [3], non_rigid_feature [P,R],
        latent [Z] -> colors [P,3]."""
        if not features.is_cuda:
            raise _gs.GsError("ColorMLPFused needs CUDA tensors (there is no CPU path)")
        L = _lib()
        dev, P = features.device, features.shape[0]
        a, keep = ColorArgs(), []
        with torch.cuda.device(dev):
            self._fill(a, features, xyz, T_fwd, camera_center, non_rigid_feature, latent, view_noise, keep)
            colors = torch.empty(P, 3, dtype=torch.float32, device=dev)
            n_in = self.feature_dim + ((self.sh_degree + 1) ** 2 - 1 if self.sh_degree > 0 else 0) + self.non_rigid_dim
            inputs = torch.empty(P, n_in, dtype=torch.float32, device=dev) if return_inputs else None
            a.colors = colors.data_ptr()
            a.inputs = inputs.data_ptr() if inputs is not None else None
            ws = self._workspace(L, a, dev)
            st = torch.cuda.current_stream(dev).cuda_stream
            if P and _saved is None:
                _gs._check(L.color_forward(ctypes.byref(a), ws.data_ptr(), ws.numel(), st), "color_forward")
            elif P:
                _gs._check(L.color_forward_train(ctypes.byref(a), ws.data_ptr(), ws.numel(), _saved.data_ptr(), _saved.numel(), st),
                           "color_forward_train")
        return (colors, inputs) if return_inputs else colors

    def saved_buffer(self, P, dev):
        return torch.empty(max(_lib().color_saved_bytes(P), 16), dtype=torch.uint8, device=dev)

    @torch.no_grad()
    def colors_backward(self, features, xyz, T_fwd, camera_center, non_rigid_feature, latent, view_noise, colors, saved, g):
        """Hand-written backward (color_backward): -> (dfeatures, dxyz, dT_fwd, dnon_rigid_feature, dlatent, [dW0, db0, ..., db2])."""
        L = _lib()
        dev, P = features.device, features.shape[0]
        b, keep = ColorBackwardArgs(), []
        with torch.cuda.device(dev):
            self._fill(b.fwd, features, xyz, T_fwd, camera_center, non_rigid_feature, latent, view_noise, keep)
            f32 = dict(dtype=torch.float32, device=dev)
            colors, g = colors.contiguous(), g.detach().contiguous().float()
            b.fwd.colors, b.dL_dcolors = colors.data_ptr(), g.data_ptr()
            dfeat = torch.empty(P, self.feature_dim, **f32)
            dxyz = torch.zeros(P, 3, **f32) if xyz is not None else None
            dT = torch.empty(P, 16, **f32) if (T_fwd is not None and self.cano_view_dir and self.sh_degree > 0) else None
            dnr = torch.empty(P, self.non_rigid_dim, **f32) if non_rigid_feature is not None else None
            dlat = torch.empty(self.latent_dim, **f32) if latent is not None else None
            dWb = [torch.empty_like(t, dtype=torch.float32) for t in self.params()]
            pn = lambda t: None if t is None else t.data_ptr()
            b.dL_dfeatures, b.dL_dxyz, b.dL_dT_fwd, b.dL_dnon_rigid_feature, b.dL_dlatent = pn(dfeat), pn(dxyz), pn(dT), pn(dnr), pn(dlat)
            (b.dL_dW0, b.dL_db0, b.dL_dW1, b.dL_db1, b.dL_dW2, b.dL_db2) = (t.data_ptr() for t in dWb)
            ws = self._workspace(L, b.fwd, dev)
            if P:
                _gs._check(L.color_backward(ctypes.byref(b), ws.data_ptr(), ws.numel(), saved.data_ptr(), saved.numel(),
                                            torch.cuda.current_stream(dev).cuda_stream), "color_backward")
            else:
                for t in dWb + ([dlat] if dlat is not None else []):
                    t.zero_()
        return dfeat, dxyz, dT, dnr, dlat, dWb

    def colors_autograd(self, features, xyz, T_fwd, camera_center, non_rigid_feature=None, latent=None, view_noise=None):
        """Differentiable `colors`: gradients to features, xyz, T_fwd, the non-rigid feature, the latent code and the three
        linear layers."""
        return _ColorFn.apply(self, view_noise, camera_center, features, xyz, T_fwd, non_rigid_feature, latent, *self.params())

    def forward(self, gaussians, camera):
        """ColorMLP.forward(gaussians, camera), texture.py:121-125: reads get_features, get_xyz, fwd_transform,
        non_rigid_feature from the deformed Gaussians and camera_center / frame_id from the camera."""
        feats = gaussians.get_features
        feats = feats.squeeze(-1) if feats.dim() == 3 else feats
        latent = self.latent_code(getattr(camera, "latent_idx", 0)) if self.latent_dim > 0 else None
        nrf = getattr(gaussians, "non_rigid_feature", None) if self.non_rigid_dim > 0 else None
        return self.colors_autograd(feats, gaussians.get_xyz, gaussians.fwd_transform, camera.camera_center, nrf, latent,
                                    getattr(camera, "view_noise", None) if self.training else None)


class _ColorFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, view_noise, camera_center, features, xyz, T_fwd, nrf, latent, *params):
        ctx.mod, ctx.view_noise, ctx.camera_center = mod, view_noise, camera_center
        ctx.none = [t is None for t in (features, xyz, T_fwd, nrf, latent)]
        saved = mod.saved_buffer(features.shape[0], features.device)
        col = mod.colors(features, xyz, T_fwd, camera_center, nrf, latent, view_noise, _saved=saved)
        ctx.save_for_backward(*[t for t in (features, xyz, T_fwd, nrf, latent) if t is not None], col, saved)
        ctx.tfwd_shape = None if T_fwd is None else T_fwd.shape
        return col

    @staticmethod
    def backward(ctx, g):
        mod = ctx.mod
        saved_t = list(ctx.saved_tensors)
        saved, col = saved_t.pop(), saved_t.pop()
        ins = [None if is_none else saved_t.pop(0) for is_none in ctx.none]
        dfeat, dxyz, dT, dnr, dlat, dWb = mod.colors_backward(ins[0], ins[1], ins[2], ctx.camera_center, ins[3], ins[4], ctx.view_noise,
                                                              col, saved, g)
        if dT is not None:
            dT = dT.view(ctx.tfwd_shape)
        if dlat is not None:
            dlat = dlat.view(ins[4].shape)
        grads = [dfeat.view(ins[0].shape), dxyz, dT, dnr, dlat] + dWb
        need = ctx.needs_input_grad[3:]
        return (None, None, None, *[gr if (n and gr is not None) else None for gr, n in zip(grads, need)])
