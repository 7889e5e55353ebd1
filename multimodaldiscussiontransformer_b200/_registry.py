"""fairseq registration that degrades to a local registry when fairseq is absent.

The reference plugs into fairseq through ``register_model`` / ``register_model_architecture`` /
``register_criterion`` / ``register_task``.  fairseq is not installable in this
image, so the decorators below call fairseq's when it imports and otherwise record the class
under the same name in ``REGISTRY`` -- the classes stay usable either way.
"""

from __future__ import annotations

REGISTRY = {"model": {}, "arch": {}, "criterion": {}, "task": {}}

try:  # pragma: no cover - fairseq is absent in the build image
    from fairseq.criterions import FairseqCriterion, register_criterion as _fs_register_criterion
    from fairseq.models import (FairseqEncoder, FairseqEncoderModel, register_model as _fs_register_model,
                                register_model_architecture as _fs_register_arch)
    from fairseq.tasks import FairseqTask, register_task as _fs_register_task

    HAVE_FAIRSEQ = True
except Exception:  # ImportError or a broken install
    import torch.nn as nn

    HAVE_FAIRSEQ = False

    class FairseqEncoder(nn.Module):
        def __init__(self, dictionary=None):
            super().__init__()
            self.dictionary = dictionary

    class FairseqEncoderModel(nn.Module):
        def __init__(self, encoder):
            super().__init__()
            self.encoder = encoder

    class FairseqCriterion(nn.Module):
        def __init__(self, task=None):
            super().__init__()
            self.task = task

    class FairseqTask:
        def __init__(self, cfg=None):
            self.cfg = cfg


def register_model(name):
    def deco(cls):
        REGISTRY["model"][name] = cls
        return _fs_register_model(name)(cls) if HAVE_FAIRSEQ else cls
    return deco


def register_model_architecture(model_name, arch_name):
    def deco(fn):
        REGISTRY["arch"][arch_name] = (model_name, fn)
        return _fs_register_arch(model_name, arch_name)(fn) if HAVE_FAIRSEQ else fn
    return deco


def register_criterion(name, dataclass=None):
    def deco(cls):
        REGISTRY["criterion"][name] = cls
        return _fs_register_criterion(name, dataclass=dataclass)(cls) if HAVE_FAIRSEQ else cls
    return deco


def register_task(name, dataclass=None):
    def deco(cls):
        REGISTRY["task"][name] = cls
        return _fs_register_task(name, dataclass=dataclass)(cls) if HAVE_FAIRSEQ else cls
    return deco
