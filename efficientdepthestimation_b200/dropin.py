"""`install()` aliases the B200 implementations under the reference's import names so that the reference's own
scripts (`from ReSIDE.models import modules, net, resnet`, `from efficientnet_pytorch import EfficientNet`,
`from ReSIDE.models.lasinger2019 import MidasNet`, `from ReSIDE import sobel`, `from ReSIDE.util import ...`) pick them
up unchanged.  Nothing is aliased until install() is called."""
import importlib
import sys

_ALIASES = {
    "ReSIDE": "efficientdepthestimation_b200.ReSIDE",
    "ReSIDE.models": "efficientdepthestimation_b200.ReSIDE.models",
    "ReSIDE.models.net": "efficientdepthestimation_b200.ReSIDE.models.net",
    "ReSIDE.models.modules": "efficientdepthestimation_b200.ReSIDE.models.modules",
    "ReSIDE.models.resnet": "efficientdepthestimation_b200.ReSIDE.models.resnet",
    "ReSIDE.models.lasinger2019": "efficientdepthestimation_b200.ReSIDE.models.lasinger2019",
    "ReSIDE.sobel": "efficientdepthestimation_b200.ReSIDE.sobel",
    "ReSIDE.util": "efficientdepthestimation_b200.ReSIDE.util",
    "ReSIDE.train": "efficientdepthestimation_b200.ReSIDE.train",
    "efficientnet_pytorch": "efficientdepthestimation_b200.efficientnet_pytorch",
}


def install(force: bool = True) -> None:
    for alias, target in _ALIASES.items():
        if force or alias not in sys.modules:
            sys.modules[alias] = importlib.import_module(target)


def uninstall() -> None:
    for alias, target in _ALIASES.items():
        m = sys.modules.get(alias)
        if m is not None and m.__name__ == target:
            del sys.modules[alias]
