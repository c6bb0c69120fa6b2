"""Decoder loading from a DeepSDF experiment directory.

Drop-in for the one function of the reference's ``deep_sdf/workspace.py`` that
the reconstruction path uses: ``config_decoder`` (`workspace.py:202-223`).
Layout read: ``<dir>/specs.json`` (keys ``NetworkArch``, ``CodeLength``,
``NetworkSpecs``) and ``<dir>/ModelParameters/<checkpoint>.pth`` holding
``{"model_state_dict": ...}`` whose keys carry the ``module.`` prefix that
``torch.nn.DataParallel`` adds.  The reference moves the module to the GPU
(`.cuda()`); here the module is only a parameter container for the packer, so
it is moved to the GPU only if one is present.
"""
from __future__ import annotations

import importlib
import json
import os

import torch

model_params_subdir = "ModelParameters"
specifications_filename = "specs.json"


def load_experiment_specifications(experiment_directory):
    path = os.path.join(experiment_directory, specifications_filename)
    if not os.path.isfile(path):
        raise Exception(
            'The experiment directory does not include specifications file "specs.json"')
    with open(path) as f:
        return json.load(f)


def config_decoder(experiment_directory, checkpoint="latest"):
    specs = load_experiment_specifications(experiment_directory)
    arch = importlib.import_module("deep_sdf." + specs["NetworkArch"])
    decoder = arch.Decoder(specs["CodeLength"], **specs["NetworkSpecs"])
    saved = torch.load(
        os.path.join(experiment_directory, model_params_subdir, checkpoint + ".pth"),
        map_location="cpu", weights_only=False)
    state = saved["model_state_dict"]
    state = {(k[len("module."):] if k.startswith("module.") else k): v
             for k, v in state.items()}
    decoder.load_state_dict(state)
    if torch.cuda.is_available():
        decoder = decoder.cuda()
    decoder.eval()
    return decoder
