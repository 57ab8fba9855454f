"""-m gpu: the reference CALLERS' sequences executed against the drop-in, after `dropin.install()`.

The reference checkout does not exist on the GPU box, so each test restates the caller's own statements (cited
line by line) on top of the aliased `ReSIDE.*` imports -- the same import lines run against the real checkout in
tests/test_dropin.py in the build container.
  (i)   Benchmark/reside_model.py:28-72 + Benchmark/benchmark.py:739-758  (ReSIDEModel: define_model ->
        load_state_dict with 'module.' stripped -> DataParallel when >1 GPU -> .to(device) -> .eval() -> normalise ->
        model(x) -> reshape -> F.interpolate)
  (ii)  ReSIDE/test.py:16-22,30-57  (DataParallel(model).cuda() + 'module.'-prefixed state_dict, forward with
        autograd ENABLED, F.interpolate, MetricsTracker.update, edge_detection)
  (iii) ReSIDE/inference_benchmark.py:110-166  (never calls .eval(): the drop-in raises and names the one-line fix)
"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PRELUDE = r'''
import os, sys, tempfile
import torch
sys.path.insert(0, os.environ["EDE_ROOT"])
from efficientdepthestimation_b200 import dropin
dropin.install()
from oracle import reside_cpu as rc, weights as W          # test infrastructure: weights + the CPU checker
tmp = tempfile.mkdtemp()
'''

SEQ_RESIDE_MODEL = PRELUDE + r'''
from ReSIDE.models.modules import E_resnet, E_densenet, E_senet, E_efficientnet     # reside_model.py:6
from ReSIDE.train import define_model                                               # reside_model.py:7
for encoder_type, variant, arch in ((E_efficientnet, "efficientnet-b0", "enb0-hu"), (E_resnet, None, "rn50-hu")):
    sd = W.make_state_dict(arch, seed=1234)
    path = os.path.join(tmp, arch + ".pth")
    # released ResNet checkpoints carry the DataParallel prefix (reside_model.py:37-38), EfficientNet ones do not
    torch.save({("module." + k if encoder_type != E_efficientnet else k): v for k, v in sd.items()}, path)
    device = torch.device("cuda")                                                   # reside_model.py:26
    model = define_model(is_resnet=encoder_type == E_resnet, is_densenet=encoder_type == E_densenet,
                         is_senet=encoder_type == E_senet, is_efficientnet=encoder_type == E_efficientnet,
                         efficientnet_variant=variant or "efficientnet-b0")         # :28-32
    state_dict = torch.load(path)                                                   # :36
    if encoder_type != E_efficientnet:
        state_dict = {f"{key.replace('module.', '')}": value for key, value in state_dict.items()}   # :38-39
    model.load_state_dict(state_dict)                                               # :44
    num_gpus = torch.cuda.device_count()
    if num_gpus > 1:
        model = torch.nn.DataParallel(model)                                        # :46-48
    model.to(device)                                                                # :50
    norm_mean = torch.Tensor([0.485, 0.456, 0.406]).reshape(1, -1, 1, 1)
    norm_stdev = torch.Tensor([0.229, 0.224, 0.225]).reshape(1, -1, 1, 1)
    model.eval()                                                                    # Benchmark/benchmark.py:740
    # two clips of two frames whose NORMALISED values are the standard-normal inputs the seeded weights were calibrated
    # on (a random-init net is only well conditioned near its BatchNorm calibration distribution: on plain uniform
    # noise even the exact fp32 mode is 1e-3 away from the oracle and every bf16 implementation > 1e-2)
    images = (W.make_input(arch, 4, seed=3) * norm_stdev + norm_mean).reshape(2, 2, 3, 228, 304)
    with torch.no_grad():                                                           # benchmark.py:745
        shape = images.shape
        C, H, Wd = shape[-3:]
        input_ = images.reshape(-1, C, H, Wd).to(device)                            # reside_model.py:61
        input_ = (input_ - norm_mean.to(device)) / norm_stdev.to(device)            # :63-64
        for rep in range(2):                                                        # DataParallel re-replicates per call
            output = model(input_)                                                  # :66
        depth = output.reshape(shape[:-3] + output.shape[-2:])                      # :69
        outputs = torch.nn.functional.interpolate(output, size=[H, Wd], mode='bilinear', align_corners=True)
    assert depth.shape == (2, 2, 114, 152) and outputs.shape == (4, 1, 228, 304) and output.device.type == "cuda"
    with torch.no_grad():
        ref = W.forward(sd, input_.cpu(), arch)
    err = rc.relative_l1(output.cpu(), ref)
    print(arch, "gpus", num_gpus, "rel_l1", err)
    assert err <= 1e-2, err
    if num_gpus > 1:                                # every replica ran on its own device with its own plan
        plans = model.module._ede_plans
        devs = sorted({k[-1] for k in plans})
        assert len(devs) >= 2, devs
        for key, entry in plans.items():
            plan = entry[0]
            assert str(plan.weights.device) == key[-1] and str(plan.arena.device) == key[-1], key
        single = define_model(is_resnet=encoder_type == E_resnet, is_efficientnet=encoder_type == E_efficientnet,
                              efficientnet_variant=variant or "efficientnet-b0")
        single.load_state_dict(state_dict)
        single = single.to(device).eval()
        with torch.no_grad():
            y1 = single(input_)
        assert rc.relative_l1(output.cpu(), y1.cpu()) <= 2e-3
print("OK")
'''

SEQ_TEST_PY = PRELUDE + r'''
from ReSIDE.models import modules, resnet, densenet, net, senet                      # test.py:7
from ReSIDE import sobel, util                                                      # test.py:9
from ReSIDE.train import define_model                                               # test.py:11
from ReSIDE.util import MetricsTracker                                              # test.py:12
arch = "rn50-hu"
sd = W.make_state_dict(arch, seed=1234)
torch.save({"module." + k: v for k, v in sd.items()}, os.path.join(tmp, "model_resnet"))
model = define_model(is_resnet=True)                                                # test.py:16
model = torch.nn.DataParallel(model).cuda()                                         # test.py:17
state_dict = torch.load(os.path.join(tmp, "model_resnet"))                          # test.py:18
model.load_state_dict(state_dict)                                                   # test.py:22
model.eval()                                                                        # test.py:31
metrics = MetricsTracker()
image = W.make_input(arch, 2, seed=0)
depth = W.make_labels(2, 228, 304, seed=1)
depth = depth.cuda(); image = image.cuda()                                          # test.py:45-46
assert torch.is_grad_enabled()                                                      # the reference runs this with autograd on
output = model(image)                                                               # test.py:48
assert not output.requires_grad
output = torch.nn.functional.interpolate(output, size=[depth.size(2), depth.size(3)], mode='bilinear',
                                         align_corners=True)                        # test.py:49-52
metrics.update(outputs=output, labels=depth)                                        # test.py:54
def edge_detection(depth):                                                          # test.py:94-102
    get_edge = sobel.Sobel().cuda()
    edge_xy = get_edge(depth)
    edge_sobel = torch.pow(edge_xy[:, 0, :, :], 2) + torch.pow(edge_xy[:, 1, :, :], 2)
    return torch.sqrt(edge_sobel)
depth_edge = edge_detection(depth); output_edge = edge_detection(output)            # test.py:56-57
with torch.no_grad():
    ref = rc.post_resize(W.forward(sd, image.cpu(), arch), (228, 304))
ref_m = rc.metrics_from_sums(rc.metric_sums(ref, depth.cpu()))
got = metrics.to_dict()
print("abs_rel", got["abs_rel"], ref_m["abs_rel"], "delta1", got["delta1"], ref_m["delta1"])
assert abs(got["abs_rel"] - ref_m["abs_rel"]) <= 2e-2 * abs(ref_m["abs_rel"])
assert abs(got["delta1"] - ref_m["delta1"]) <= 2e-2
ref_edge = rc.edge_magnitude(ref)
assert output_edge.shape == (2, 228, 304)
assert rc.relative_l1(output_edge.cpu().reshape(2, 1, 228, 304), ref_edge.reshape(2, 1, 228, 304)) <= 5e-2
print("OK")
'''

SEQ_INFERENCE_BENCHMARK = PRELUDE + r'''
from ReSIDE.models.lasinger2019 import MidasNet                                     # inference_benchmark.py:18
from ReSIDE.train import define_model                                               # :19
from ReSIDE.util import Timer                                                       # :20
arch = "enb0-hu"
sd = W.make_state_dict(arch, seed=1234)
model_path = os.path.join(tmp, "efficientnet-b0-hu.pth")
torch.save({"module." + k: v for k, v in sd.items()}, model_path)                   # a 'module.'-prefixed checkpoint
loading_timer = Timer()
with loading_timer:
    model = define_model(is_efficientnet=True, efficientnet_variant="efficientnet-b0")   # :133
    try:
        state_dict = torch.load(model_path)
        model.load_state_dict(state_dict)                                           # :138-140: must raise RuntimeError
        raise AssertionError("a 'module.'-prefixed state_dict must not load strictly")
    except RuntimeError:
        state_dict = torch.load(model_path)
        state_dict = {key.replace("module.", ""): value for key, value in state_dict.items()}   # :144-145
        model.load_state_dict(state_dict)                                           # :147
    model.cuda()                                                                    # :150
batch = W.make_input(arch, 8, seed=0)                                               # DataLoader(batch_size=8), :55
inference_timer = Timer()
with torch.no_grad(), inference_timer:
    try:
        depth = model(batch.cuda())                                                 # :161 -- the model is still in train()
    except RuntimeError as e:
        msg = str(e)
        assert "call .eval() first" in msg and "inference forward only" in msg, msg
        print("raised:", msg)
    else:
        raise AssertionError("training-mode forward must raise (no silent running-statistics forward)")
    model.eval()                                                                    # the one line INTEGRATION.md asks for
    height, width = batch.shape[-2:]
    depth = model(batch.cuda())
    _ = torch.nn.functional.interpolate(depth, size=(height, width), mode='bilinear', align_corners=True)   # :163
peak = torch.cuda.max_memory_allocated()                                            # :165 stays meaningful
assert depth.shape == (8, 1, 114, 152) and peak > 0 and inference_timer.elapsed.total_seconds() > 0
with torch.no_grad():
    ref = W.forward(sd, batch, arch)
assert rc.relative_l1(depth.cpu(), ref) <= 1e-2
print("OK")
'''


def _run(code):
    env = dict(os.environ, EDE_ROOT=ROOT, OMP_NUM_THREADS="8")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd="/tmp", env=env, timeout=1200)
    assert r.returncode == 0 and "OK" in r.stdout, r.stdout[-3000:] + r.stderr[-4000:]
    print(r.stdout[-1500:])


def test_reside_model_sequence(cuda, lib):
    _run(SEQ_RESIDE_MODEL)


def test_test_py_sequence_dataparallel_autograd_on(cuda, lib):
    _run(SEQ_TEST_PY)


def test_inference_benchmark_sequence_without_eval(cuda, lib):
    _run(SEQ_INFERENCE_BENCHMARK)
