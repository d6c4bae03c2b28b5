"""oracle/ref_field.py — TEST INFRASTRUCTURE ONLY; works ONLY where /root/reference exists (the build
container).  Nothing that runs on the GPU box may import this module.

Imports the reference's OWN ``src/model/model.py`` under a stub meta-path finder for the third-party
packages that are absent here (pytorch3d, trimesh, matplotlib, tabulate, imageio, ffmpeg) and with
torchvision's ``vgg16`` patched to ``weights=None`` (``src/model/vgg_net.py:10`` would otherwise try to
download weights at import).  ``NeuralDisplacementField.forward`` (model.py:393-453) and
``FourierFeatureTransform`` (fourier_feature_transform.py:7-55) then run unmodified on CPU with real
autograd; ``get_meshes`` cannot (it needs real PyTorch3D ``Meshes`` / ``Transform3d``).
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest import mock

REFERENCE_ROOT = "/root/reference"
_STUBBED = ("pytorch3d", "trimesh", "matplotlib", "tabulate", "imageio", "ffmpeg")


class _Stub(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        m = mock.MagicMock(name=f"{self.__name__}.{name}")
        setattr(self, name, m)
        return m


class _Finder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, name, path, target=None):
        if name.split(".")[0] in _STUBBED:
            return importlib.machinery.ModuleSpec(name, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _Stub(spec.name)
        m.__path__ = []
        m.__all__ = []
        return m

    def exec_module(self, module):
        pass


_ref_module = None


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "model"))


def import_reference_model():
    """-> the reference's ``src.model.model`` module (cached)."""
    global _ref_module
    if _ref_module is not None:
        return _ref_module
    if not available():
        raise RuntimeError("reference tree not present: oracle/ref_field.py only works in the build container")
    sys.meta_path.insert(0, _Finder())
    import torchvision.models as tvm
    import torchvision.models.vgg as vgg
    orig = vgg.vgg16
    patched = lambda *a, **k: orig(weights=None)
    vgg.vgg16 = patched
    tvm.vgg16 = patched
    for p in (REFERENCE_ROOT, os.path.join(REFERENCE_ROOT, "src"), os.path.join(REFERENCE_ROOT, "src", "model")):
        if p not in sys.path:
            sys.path.insert(0, p)
    argv, sys.argv = sys.argv, ["x"]
    try:
        import src.model.model as ref_model   # noqa: E402  (the reference's own module)
    finally:
        sys.argv = argv
    _ref_module = ref_model
    return ref_model


def make_reference_field(train_size: int = 8, val_size: int = 2, **kw):
    """The reference's NeuralDisplacementField with train.py:148-157's configuration, on CPU."""
    M = import_reference_model()
    args = dict(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=train_size, val_size=val_size,
                shapevec_size=100, texvec_size=100, posevec_size=100, device="cpu", opts=None)
    args.update(kw)
    return M.NeuralDisplacementField(**args)
