"""Import the UNMODIFIED reference (`/root/reference`) in the build container.
TEST INFRASTRUCTURE - used only by `oracle/make_goldens.py` (golden-vector
generation) here; `/root/reference` does not exist on the GPU box.

Shims (harness side only, reference files untouched - SURVEY.md Appendix B):
  * stub modules for `addict`, `plyfile`, `skimage.measure` (absent here);
  * on a CPU-only host: identity `.cuda()` and no-op `torch.cuda.synchronize`.
"""
from __future__ import annotations

import sys
import types

REF_ROOT = "/root/reference"


class _Dict(dict):
    """Stand-in for addict.Dict: attribute access, recursive wrapping, __missing__ hook."""

    def __init__(self, *a, **kw):
        super().__init__()
        for k, v in dict(*a, **kw).items():
            self[k] = v

    def __setitem__(self, k, v):
        if isinstance(v, dict) and not isinstance(v, _Dict):
            v = type(self)(**v)
        super().__setitem__(k, v)

    def __getattr__(self, k):
        if k.startswith("__"):
            raise AttributeError(k)
        try:
            return self[k]
        except KeyError:
            return self.__missing__(k)

    def __missing__(self, k):
        v = type(self)()
        self[k] = v
        return v

    def __setattr__(self, k, v):
        self[k] = v


def import_reference():
    """Returns the reference's modules as a namespace; must run in a process
    that has NOT imported this repo's own `reconstruct` / `deep_sdf` mirrors."""
    import torch
    for name in ("reconstruct", "deep_sdf"):
        if name in sys.modules and REF_ROOT not in getattr(sys.modules[name], "__file__", ""):
            raise RuntimeError("repo mirror of %s already imported in this process" % name)
    addict = types.ModuleType("addict")
    addict.Dict = _Dict
    sys.modules.setdefault("addict", addict)
    sys.modules.setdefault("plyfile", types.ModuleType("plyfile"))
    sk = types.ModuleType("skimage")
    skm = types.ModuleType("skimage.measure")
    sk.measure = skm
    sys.modules.setdefault("skimage", sk)
    sys.modules.setdefault("skimage.measure", skm)
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
        torch.nn.Module.cuda = lambda self, *a, **k: self
        torch.cuda.synchronize = lambda *a, **k: None
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import reconstruct.optimizer as ropt
    import reconstruct.loss as rloss
    import reconstruct.loss_utils as rlu
    import reconstruct.utils as rutils
    import deep_sdf.deep_sdf_decoder as rdec
    import deep_sdf.workspace as rws
    return types.SimpleNamespace(optimizer=ropt, loss=rloss, loss_utils=rlu, utils=rutils,
                                 decoder=rdec, workspace=rws)
