"""Import the UNMODIFIED reference: `/root/reference` in the build container, else the verbatim copy of its
hot-path files under `oracle/_ref/` (made by `oracle/fetch_ref.py`, git-ignored, travels to the GPU box).
TEST INFRASTRUCTURE - used by `oracle/make_goldens*.py` (golden-vector generation), `oracle/ref_runner.py`
(the timed reference arm of bench.py) and tests; never by the product.

Shims (harness side only, reference files untouched - SURVEY.md Appendix B):
  * stub modules for `addict`, `plyfile`, `skimage.measure` (absent here);
  * on a CPU-only host: identity `.cuda()` and no-op `torch.cuda.synchronize`.
"""
from __future__ import annotations

import os
import sys
import types

_LOCAL = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
REF_ROOT = os.environ.get("DSPSLAM_REF_ROOT") or ("/root/reference" if os.path.isdir("/root/reference/reconstruct") else _LOCAL)


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "reconstruct", "optimizer.py"))


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


def import_reference(force_cpu=False):
    """Returns the reference's modules as a namespace; must run in a process
    that has NOT imported this repo's own `reconstruct` / `deep_sdf` mirrors.
    force_cpu: make the reference's unconditional `.cuda()` calls no-ops even on a host with a GPU
    (the CPU-timed arm); otherwise they are left alone where CUDA is available."""
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
    if force_cpu or not torch.cuda.is_available():
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
