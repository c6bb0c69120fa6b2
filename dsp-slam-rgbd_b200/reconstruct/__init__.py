"""Drop-in for the reference's ``reconstruct`` package on B200 (`reconstruct/__init__.py:1-23`).

`src/System.cc:99` does ``py::module::import("reconstruct").attr("get_sequence")(strSequencePath, pyCfg)``
right after ``reconstruct.utils.get_configs / get_decoder`` (`:95-98`); both names live here with the
reference's signatures.

Modules of the reference that are NOT on the reconstruction hot path and therefore not mirrored
(`detector2d`, `detector3d`, `loss`, `loss_utils`, ...) still resolve: when another ``reconstruct/`` directory
(the reference checkout, which `src/System.cc:94` appends to ``sys.path`` as ``./``) is found further down
``sys.path``, it is appended to this package's ``__path__``; mirrored modules keep precedence.
"""
import os
import sys


def _extend_path():
    here = os.path.dirname(os.path.abspath(__file__))
    for entry in list(sys.path):
        cand = os.path.abspath(os.path.join(entry or ".", "reconstruct"))
        if cand != here and os.path.isfile(os.path.join(cand, "__init__.py")) and cand not in __path__:
            __path__.append(cand)


_extend_path()


def get_detectors(configs):
    """`reconstruct/__init__.py:1-13`.  Off-line (pre-stored labels): ``(None, None)`` for KITTI, ``None``
    otherwise.  On-line detectors (mmdetection / mmdetection3d) are outside this repo's scope; they load from the
    reference checkout if it is on ``sys.path`` (see module docstring)."""
    if configs.detect_online:
        _extend_path()
        try:
            from .detector2d import get_detector2d
            if configs.data_type == "KITTI":
                from .detector3d import get_detector3d
                return get_detector2d(configs), get_detector3d(configs)
            return get_detector2d(configs)
        except ImportError as e:
            raise ImportError("detect_online needs the reference's reconstruct/detector2d.py / detector3d.py "
                              "(mmdetection) on sys.path; the B200 drop-in covers pre-stored labels only") from e
    if configs.data_type == "KITTI":
        return None, None
    return None


def get_sequence(data_dir, configs):
    """`reconstruct/__init__.py:16-23`: KITTI -> KITIISequence, Redwood / Freiburg -> MonoSequence
    (anything else returns None, as the reference does)."""
    if configs.data_type == "KITTI":
        from .kitti_sequence import KITIISequence
        return KITIISequence(data_dir, configs)
    if configs.data_type == "Redwood" or configs.data_type == "Freiburg":
        from .mono_sequence import MonoSequence
        return MonoSequence(data_dir, configs)
