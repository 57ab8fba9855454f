"""`install()` puts the B200 implementations under the reference's import names so that the reference's own
scripts (`from ReSIDE.models import modules, resnet, densenet, net, senet`, `from ReSIDE.train import define_model`,
`from ReSIDE.models.lasinger2019 import MidasNet`, `from ReSIDE import sobel, loaddata, util`, ...) pick them up
unchanged.  Nothing is aliased until install() is called.

Two situations:

* **The reference checkout is importable** (a maintainer runs its scripts from the repository root, so `ReSIDE` is a
  real package on `sys.path`).  Only the LEAF modules that this package re-implements are replaced --
  `ReSIDE.models.{net,modules,resnet,lasinger2019}` and `ReSIDE.sobel` -- everything else of the reference
  (`ReSIDE.train`, `ReSIDE.loaddata`, `ReSIDE.demo_transform`, `ReSIDE.models.{densenet,senet}`, `ReSIDE.util`, ...)
  stays the reference's own code: its `define_model` (`ReSIDE/train.py:20-38`) then builds the B200 modules because it
  looks the classes up through the replaced leaf modules.  `ReSIDE.util.MetricsTracker` is swapped for the fused
  one (same interface, `patch_metrics=False` keeps the reference's).
* **No reference checkout** (tests on the GPU box, stand-alone use): the mirror package under
  `efficientdepthestimation_b200.ReSIDE` is aliased as `ReSIDE` as a whole; it carries stub `densenet` / `senet`
  modules and `E_densenet` / `E_senet` classes that raise on construction, so the reference's import lines work.
"""
import importlib
import importlib.util
import sys

_PKG = "efficientdepthestimation_b200"
# leaf modules re-implemented on the B200 path (replaced in both situations)
_LEAVES = ("ReSIDE.models.net", "ReSIDE.models.modules", "ReSIDE.models.resnet", "ReSIDE.models.lasinger2019",
           "ReSIDE.sobel")
# the rest of the mirror: only aliased when there is no real `ReSIDE` package to keep
_MIRROR_ONLY = ("ReSIDE", "ReSIDE.models", "ReSIDE.models.densenet", "ReSIDE.models.senet", "ReSIDE.util",
                "ReSIDE.train")
_installed = {}          # alias -> previous sys.modules entry (None = absent), for uninstall()
_patched = []            # (module, attribute, previous value)


def _is_ours(mod) -> bool:
    return mod is not None and getattr(mod, "__name__", "").startswith(_PKG + ".")


def _real_reside():
    """The reference's own `ReSIDE` package if one is importable, else None (never imports submodules)."""
    cur = sys.modules.get("ReSIDE")
    if cur is not None:
        return None if _is_ours(cur) else cur
    try:
        spec = importlib.util.find_spec("ReSIDE")
    except (ImportError, ValueError):
        spec = None
    if spec is None:
        return None
    return importlib.import_module("ReSIDE")


def _alias(alias: str, force: bool) -> None:
    if not force and alias in sys.modules:
        return
    target = importlib.import_module(f"{_PKG}.{alias}")
    _installed.setdefault(alias, sys.modules.get(alias))
    sys.modules[alias] = target
    parent, _, leaf = alias.rpartition(".")
    if parent and parent in sys.modules:                  # `from ReSIDE.models import modules` reads the attribute first
        p = sys.modules[parent]
        if not _is_ours(p):
            _patched.append((p, leaf, getattr(p, leaf, None)))
        setattr(p, leaf, target)


def install(force: bool = True, patch_metrics: bool = True) -> str:
    """Returns 'leaves' (a real reference package was found and kept) or 'mirror' (whole mirror package aliased)."""
    real = _real_reside()
    ep = "efficientnet_pytorch"
    if force or ep not in sys.modules:
        _installed.setdefault(ep, sys.modules.get(ep))
        sys.modules[ep] = importlib.import_module(f"{_PKG}.{ep}")
    if real is None:
        for alias in _MIRROR_ONLY + _LEAVES:
            _alias(alias, force)
        return "mirror"
    importlib.import_module("ReSIDE.models")              # the real package (an empty __init__ in the reference)
    for alias in _LEAVES:
        _alias(alias, force)
    if patch_metrics:
        util = importlib.import_module("ReSIDE.util")
        ours = importlib.import_module(f"{_PKG}.ReSIDE.util")
        _patched.append((util, "MetricsTracker", getattr(util, "MetricsTracker", None)))
        util.MetricsTracker = ours.MetricsTracker
        for name in ("ReSIDE.train", "ReSIDE.test"):      # `from ReSIDE.util import MetricsTracker` already executed
            m = sys.modules.get(name)
            if m is not None and hasattr(m, "MetricsTracker"):
                _patched.append((m, "MetricsTracker", m.MetricsTracker))
                m.MetricsTracker = ours.MetricsTracker
    return "leaves"


def uninstall() -> None:
    for mod, attr, prev in reversed(_patched):
        if prev is None:
            if hasattr(mod, attr):
                delattr(mod, attr)
        else:
            setattr(mod, attr, prev)
    _patched.clear()
    for alias, prev in _installed.items():
        if prev is None:
            sys.modules.pop(alias, None)
        else:
            sys.modules[alias] = prev
    _installed.clear()
