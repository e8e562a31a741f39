"""TEST INFRASTRUCTURE ONLY -- live-reference harness (never imported by the product path).

Runs the UNMODIFIED reference scripts from /root/reference (present only in the build
container, never on the GPU box) so that golden vectors can be generated from them:

* ``import_reference()``  -> the reference modules imported as libraries (their work is
  all under ``if __name__ == "__main__"`` so importing is side-effect free) for
  per-function goldens (tools/projection_2d_to_3d.py:37-301, tools/refinement.py:26-90).
* ``run_cli(script, config, cls, cwd)`` -> ``runpy.run_path(..., run_name="__main__")``
  for end-to-end goldens (tools/projection_2d_to_3d.py:336-634, tools/refinement.py:135-428).

Modules that are not installed here and do not touch the arithmetic are stubbed
(SURVEY.md section 8c): matplotlib, configs.config, munch, open3d, clip.
"""
from __future__ import annotations

import contextlib
import hashlib
import importlib
import os
import runpy
import sys
import types

REFERENCE_ROOT = os.environ.get("BFF_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "tools", "projection_2d_to_3d.py"))


class _Munch(dict):
    """Attribute dict standing in for munch.Munch (projection_2d_to_3d.py:14,339)."""

    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__

    @classmethod
    def fromDict(cls, d):
        return cls(d)


def fake_clip_similarity_vector(text: str, dim: int = 32):
    """Deterministic text embedding used by the fake ``clip`` on BOTH sides of a parity
    test: CLIP is an input to the path (refinement.py:93-117), not part of it."""
    import torch

    seed = int.from_bytes(hashlib.sha1(text.encode()).digest()[:4], "little")
    g = torch.Generator().manual_seed(seed)
    return torch.randn(1, dim, generator=g) + 1.0


def _install_stubs():
    import torch

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    if "matplotlib" not in sys.modules:
        try:
            importlib.import_module("matplotlib.pyplot")
        except Exception:
            mp = mod("matplotlib")
            mp.pyplot = mod("matplotlib.pyplot")
    cfgmod = mod("configs.config", min_aggragated_masks=2)
    pkg = mod("configs", config=cfgmod)
    pkg.__path__ = []
    mod("munch", Munch=_Munch)
    if "open3d" not in sys.modules:
        mod("open3d")

    class _Tok:
        def __init__(self, texts):
            self.texts = texts

        def to(self, *_a, **_k):
            return self

    class _Model:
        def encode_text(self, tok):
            return torch.cat([fake_clip_similarity_vector(t) for t in tok.texts], dim=0)

    mod("clip", load=lambda name, device=None: (_Model(), None), tokenize=lambda texts: _Tok(list(texts)))


@contextlib.contextmanager
def reference_env(cwd: str | None = None):
    """sys.path / cwd / env needed by the reference."""
    old_path = list(sys.path)
    old_cwd = os.getcwd()
    old_env = {k: os.environ.get(k) for k in ("TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD", "CUDA_VISIBLE_DEVICES")}
    os.environ["TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD"] = "1"
    os.environ["CUDA_VISIBLE_DEVICES"] = ""
    _install_stubs()
    sys.path.insert(0, os.path.join(REFERENCE_ROOT, "tools"))
    sys.path.insert(0, REFERENCE_ROOT)  # for `evaluation.dataset.scannet200`
    if cwd is not None:
        os.makedirs(os.path.join(cwd, "checkpoints"), exist_ok=True)
        os.chdir(cwd)
    try:
        yield
    finally:
        os.chdir(old_cwd)
        sys.path[:] = old_path
        for k, v in old_env.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


_cached = {}


def import_reference():
    """Return (projection_module, refinement_module, rle_module) of the real reference."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    if "mods" not in _cached:
        with reference_env():
            P = importlib.import_module("projection_2d_to_3d")
            R = importlib.import_module("refinement")
            E = importlib.import_module("utils.rle_encode_decode")
        _cached["mods"] = (P, R, E)
    return _cached["mods"]


def run_cli(script: str, config_path: str, cls: str, cwd: str):
    """Run a reference CLI (``projection_2d_to_3d.py`` or ``refinement.py``) in-process."""
    path = os.path.join(REFERENCE_ROOT, "tools", script)
    old_argv = list(sys.argv)
    sys.argv = [path, "--config", config_path, "--cls", cls]
    try:
        with reference_env(cwd):
            runpy.run_path(path, run_name="__main__")
    finally:
        sys.argv = old_argv
