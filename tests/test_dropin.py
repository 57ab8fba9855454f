"""not gpu: `dropin.install()` against (a) the real reference checkout (when /root/reference exists -- this container),
(b) a minimal fake reference tree (runs everywhere) and (c) no reference at all (mirror mode, the GPU box).
Each case executes the IMPORT LINES of the reference's callers (ReSIDE/test.py:7-12, test_nyu.py:14-16,
inference_benchmark.py:17-20, Benchmark/reside_model.py:6-7, train.py:14-17) in a fresh interpreter."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CALLER_IMPORTS = """
from ReSIDE.models import modules, resnet, densenet, net, senet          # test.py:7, train.py:15, demo.py:10
from ReSIDE import sobel, util                                           # test.py:9
from ReSIDE.models.lasinger2019 import MidasNet, Encoder, Decoder        # test.py:10, train.py:16
from ReSIDE.train import define_model                                    # test.py:11, test_nyu.py:16
from ReSIDE.util import MetricsTracker, Timer                            # test.py:12, inference_benchmark.py:20
from ReSIDE.models.modules import E_resnet, E_densenet, E_senet, E_efficientnet   # Benchmark/reside_model.py:6
"""

CHECKS = """
P = "efficientdepthestimation_b200."
assert modules.__name__.startswith(P) and net.__name__.startswith(P) and resnet.__name__.startswith(P)
assert sobel.__name__.startswith(P) and MidasNet.__module__.startswith(P)
assert MetricsTracker.__module__.startswith(P)
m = define_model(is_resnet=True)
assert type(m).__module__.startswith(P) and type(m.E).__module__.startswith(P), type(m)
m = define_model(is_efficientnet=True, efficientnet_variant="efficientnet-b0")
assert type(m.E).__module__.startswith(P) and len(m.state_dict()) == 522
# the encoder-type comparisons of Benchmark/reside_model.py:28-32 work, the unsupported encoders raise RuntimeError
assert E_resnet is modules.E_resnet and E_efficientnet is modules.E_efficientnet
for bad in (E_densenet, E_senet):
    try:
        bad(None)
    except RuntimeError as e:
        assert "Unsupported encoder" in str(e)
    else:
        raise AssertionError("constructing an unsupported encoder must raise")
from efficientdepthestimation_b200 import dropin
dropin.uninstall()
import sys
assert not any(getattr(v, "__name__", "").startswith(P) for k, v in sys.modules.items() if k.startswith("ReSIDE"))
print("OK", MODE)
"""


def run(code, extra_path=()):
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([*extra_path, ROOT]), OMP_NUM_THREADS="2", WANDB_MODE="disabled")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd="/tmp", env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r.stdout


def test_mirror_mode_without_a_reference_checkout():
    out = run("from efficientdepthestimation_b200 import dropin\nMODE = dropin.install()\nassert MODE == 'mirror'\n"
              + CALLER_IMPORTS + CHECKS)
    assert "OK mirror" in out


def test_leaf_mode_against_a_fake_reference_tree(tmp_path):
    """A stand-in checkout with the modules the B200 package does NOT re-implement: they must stay the checkout's."""
    pk = tmp_path / "ReSIDE"
    (pk / "models").mkdir(parents=True)
    (pk / "__init__.py").write_text("")
    (pk / "models" / "__init__.py").write_text("")
    for name in ("densenet", "senet", "net", "modules", "resnet", "lasinger2019"):
        (pk / "models" / f"{name}.py").write_text(f"ORIGIN = 'checkout'\ndef {name}161(pretrained=False):\n    return None\n")
    (pk / "sobel.py").write_text("ORIGIN = 'checkout'\n")
    (pk / "loaddata.py").write_text("ORIGIN = 'checkout'\n")
    (pk / "demo_transform.py").write_text("class Scale: pass\nclass CenterCrop: pass\nclass ToTensor: pass\nclass Normalize: pass\n")
    (pk / "util.py").write_text("class MetricsTracker:\n    ORIGIN = 'checkout'\nclass Timer:\n    ORIGIN = 'checkout'\n")
    (pk / "train.py").write_text(textwrap.dedent("""
        from ReSIDE import sobel, loaddata
        from ReSIDE.models import modules, resnet, densenet, net, senet
        from ReSIDE.models.lasinger2019 import MidasNet, Encoder, Decoder
        from ReSIDE.util import MetricsTracker, Timer
        def define_model(is_resnet=False, is_densenet=False, is_senet=False, is_efficientnet=False,
                         efficientnet_variant='efficientnet-b0'):
            if is_resnet:
                return net.model(modules.E_resnet(resnet.resnet50(pretrained=True)), num_features=2048,
                                 block_channel=[256, 512, 1024, 2048])
            enc = modules.E_efficientnet(name=efficientnet_variant, pretrained=True)
            return net.model(enc, num_features=enc.num_features, block_channel=enc.block_out_channels)
        """))
    code = ("from efficientdepthestimation_b200 import dropin\nMODE = dropin.install()\nassert MODE == 'leaves'\n"
            + CALLER_IMPORTS +
            "from ReSIDE import loaddata\nfrom ReSIDE.demo_transform import Scale, CenterCrop, ToTensor, Normalize\n"
            "import ReSIDE.train as T\n"
            "assert loaddata.ORIGIN == 'checkout' and densenet.ORIGIN == 'checkout' and senet.ORIGIN == 'checkout'\n"
            "assert Timer.ORIGIN == 'checkout' and T.__file__.startswith(%r)\n" % str(tmp_path) + CHECKS)
    assert "OK leaves" in run(code, extra_path=[str(tmp_path)])


@pytest.mark.skipif(not os.path.isdir("/root/reference/ReSIDE"), reason="reference checkout not present on this machine")
def test_leaf_mode_against_the_real_reference_checkout():
    """The reference's OWN ReSIDE/train.py:define_model builds the B200 modules after install()."""
    code = ("from efficientdepthestimation_b200 import dropin\nMODE = dropin.install()\nassert MODE == 'leaves'\n"
            + CALLER_IMPORTS +
            "from ReSIDE import loaddata                                     # test.py:9, test_nyu.py:14\n"
            "from ReSIDE.demo_transform import Scale, CenterCrop, ToTensor, Normalize   # inference_benchmark.py:17\n"
            "import ReSIDE.train as T\n"
            "assert T.__file__ == '/root/reference/ReSIDE/train.py', T.__file__\n"
            "assert densenet.__file__.startswith('/root/reference/') and loaddata.__file__.startswith('/root/reference/')\n"
            + CHECKS)
    assert "OK leaves" in run(code, extra_path=["/root/reference"])
