"""Decoder parameters -> effective FP32 weights for the CUDA library.

Weight norm is folded once at load (`w = g * v / ||v||`, norm per output row: torch
`weight_norm` default dim=0, reference `deep_sdf/deep_sdf_decoder.py:49-54`), where the
reference recomputes it in a forward pre-hook on every call.  The device-side packing (transposed
FP32 copies, FP16 tcgen05 operand tiles) happens inside the library (csrc/api.cu, kernels_tc.cu).
"""
from __future__ import annotations

import numpy as np


def _np(v):
    return v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v)


def effective_weights(state_dict):
    """[(W_k float32 [out,in], b_k float32 [out])] for lin0..linK from a DeepSDF state_dict
    (old-style `weight_g/weight_v`, parametrization-style `original0/1`, or plain `weight`)."""
    sd = {(k[7:] if k.startswith("module.") else k): v for k, v in state_dict.items()}
    layers = []
    k = 0
    while ("lin%d.bias" % k) in sd:
        p = "lin%d." % k
        if p + "weight_g" in sd:
            g, v = _np(sd[p + "weight_g"]), _np(sd[p + "weight_v"])
        elif p + "parametrizations.weight.original0" in sd:
            g, v = _np(sd[p + "parametrizations.weight.original0"]), _np(sd[p + "parametrizations.weight.original1"])
        else:
            g, v = None, _np(sd[p + "weight"])
        v = v.astype(np.float64)
        if g is not None:
            v = v * (g.astype(np.float64).reshape(-1, 1) / np.linalg.norm(v, axis=1, keepdims=True))
        layers.append((np.ascontiguousarray(v.astype(np.float32)),
                       np.ascontiguousarray(_np(sd[p + "bias"]).astype(np.float32))))
        k += 1
    return layers


def check_variant(decoder, state_dict):
    """Raise ValueError for DeepSDF variants whose arithmetic the library does not implement (evaluating them as the
    plain net would give silently wrong SDFs): `use_tanh` (the reference applies tanh twice then,
    deep_sdf_decoder.py:97-110), LayerNorm layers (`bnK.*` parameters: weight_norm=False with norm_layers), an extra
    xyz re-injection (`xyz_in_all`), active dropout is irrelevant in eval mode."""
    if getattr(decoder, "use_tanh", False):
        raise ValueError("unsupported decoder variant: use_tanh=True (double tanh) is not implemented")
    if getattr(decoder, "xyz_in_all", False):
        raise ValueError("unsupported decoder variant: xyz_in_all=True is not implemented")
    bad = [k for k in state_dict if k.split(".")[-2 if "." in k else 0].startswith("bn") or k.startswith("bn") or ".bn" in k]
    if bad:
        raise ValueError("unsupported decoder variant: LayerNorm parameters %s (weight_norm=False with norm_layers)" % bad[:3])
    li = getattr(decoder, "latent_in", None)
    if li is not None and tuple(li) != (4,):
        raise ValueError("unsupported decoder variant: latent_in=%s (supported: [4])" % (list(li),))


def check_supported(layers):
    """Raise ValueError unless the layer shapes are the 8x512 latent_in=[4] family; returns code_len."""
    if len(layers) != 9:
        raise ValueError("decoder must have 9 linear layers (8 hidden + output), got %d" % len(layers))
    in_dim = layers[0][0].shape[1]
    L = in_dim - 3
    want = [(512, in_dim), (512, 512), (512, 512), (512 - in_dim, 512), (512, 512), (512, 512), (512, 512),
            (512, 512), (1, 512)]
    got = [w.shape for w, _ in layers]
    if got != want or not (1 <= L <= 64):
        raise ValueError("unsupported decoder geometry %s (supported: 8x512, latent_in=[4], code_len<=64)" % (got,))
    return L
