"""Generate the golden fixtures that pin the CPU oracle against the REAL reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py

For every architecture it (1) builds the oracle's seeded state_dict, (2) loads it STRICTLY
into the reference's own nn.Modules (ReSIDE/models/net.py, modules.py, resnet.py,
lasinger2019.py imported from /root/reference on top of oracle/_shim), (3) runs the
reference forward on the seeded input and stores the output (+ encoder/D/MFF checksums),
(4) asserts that oracle.reside_cpu reproduces it.  Sobel / loss / metric fixtures come from
ReSIDE/sobel.py, the verbatim loss lines of ReSIDE/train.py:216-236 and ReSIDE/util.py's
MetricsTracker.  Output: tests/golden/*.pt (small fp32 tensors).
"""
import os
import sys

import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_shim"), "/root/reference", ROOT]

from oracle import reside_cpu as rc          # noqa: E402
from oracle import weights as W              # noqa: E402

from ReSIDE import sobel as ref_sobel                                   # noqa: E402
from ReSIDE.models import modules as ref_modules, net as ref_net, resnet as ref_resnet  # noqa: E402
from ReSIDE.models.lasinger2019 import Decoder, Encoder, MidasNet       # noqa: E402
from ReSIDE.util import MetricsTracker                                  # noqa: E402


def build_reference(arch, lr_out_wh=None):
    enc, dec = arch.split("-")
    name = W.encoder_name(arch)
    if dec == "hu":
        if enc == "rn50":
            return ref_net.model(ref_modules.E_resnet(ref_resnet.resnet50(pretrained=False)),
                                 2048, [256, 512, 1024, 2048])
        E = ref_modules.E_efficientnet(name, pretrained=False)
        return ref_net.model(E, E.num_features, E.block_out_channels)
    e = Encoder(name, pretrained=False)
    return MidasNet(e, Decoder(e.block_out_channels, num_features="auto"), output_size=lr_out_wh,
                    input_size=tuple(reversed(W.input_hw(arch))))


def reference_loss(output, depth):
    """ReSIDE/train.py:202,216-236 with `.cuda()` dropped (CPU tensors)."""
    cos = nn.CosineSimilarity(dim=1, eps=0)
    get_gradient = ref_sobel.Sobel()
    ones = torch.ones(depth.size(0), 1, depth.size(2), depth.size(3)).float()
    depth_grad = get_gradient(depth)
    output_grad = get_gradient(output)
    depth_grad_dx = depth_grad[:, 0, :, :].contiguous().view_as(depth)
    depth_grad_dy = depth_grad[:, 1, :, :].contiguous().view_as(depth)
    output_grad_dx = output_grad[:, 0, :, :].contiguous().view_as(depth)
    output_grad_dy = output_grad[:, 1, :, :].contiguous().view_as(depth)
    depth_normal = torch.cat((-depth_grad_dx, -depth_grad_dy, ones), 1)
    output_normal = torch.cat((-output_grad_dx, -output_grad_dy, ones), 1)
    loss_depth = torch.log(torch.abs(output - depth) + 0.5).mean()
    loss_dx = torch.log(torch.abs(output_grad_dx - depth_grad_dx) + 0.5).mean()
    loss_dy = torch.log(torch.abs(output_grad_dy - depth_grad_dy) + 0.5).mean()
    loss_normal = torch.abs(1 - cos(output_normal, depth_normal)).mean()
    return {"depth": loss_depth, "dx": loss_dx, "dy": loss_dy, "normal": loss_normal,
            "loss": loss_depth + loss_normal + (loss_dx + loss_dy)}


def main():
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    report = {}
    for arch in W.ARCHS:
        sd = W.make_state_dict(arch, seed=1234)
        hw = W.input_hw(arch)
        lr_out_hw = (hw[0] // 2, hw[1] // 2)
        ref = build_reference(arch, lr_out_wh=(lr_out_hw[1], lr_out_hw[0]))
        ref.load_state_dict(sd, strict=True)                      # pins the key layout
        assert list(ref.state_dict().keys()) == list(sd.keys()), arch
        ref.eval()
        x = W.make_input(arch, 1, seed=0)
        with torch.no_grad():
            y_ref = ref(x)
            taps = {}
            y_or = W.forward(sd, x, arch, lr_output_hw=lr_out_hw, taps=taps)
        err = float((y_ref - y_or).abs().max())
        assert err <= 1e-5 * float(y_ref.abs().max()), (arch, err)
        fixture = {"arch": arch, "seed": 1234, "input_seed": 0, "output": y_ref.clone(),
                   "n_params": sum(v.numel() for k, v in sd.items()
                                   if not k.endswith(("running_mean", "running_var", "num_batches_tracked"))),
                   "n_entries": len(sd),
                   "tap_abs_mean": {k: float(v.abs().mean()) for k, v in taps.items()},
                   "tap_shapes": {k: tuple(v.shape) for k, v in taps.items()}}
        torch.save(fixture, os.path.join(HERE, f"{arch}.pt"))
        report[arch] = (tuple(y_ref.shape), float(y_ref.min()), float(y_ref.mean()), float(y_ref.max()), err)
        print(arch, report[arch], flush=True)

    # Sobel / loss / metrics fixtures (reference: sobel.py, train.py:216-236, util.py:39-92)
    out = 0.5 + 9.5 * torch.rand(3, 1, 57, 76, generator=torch.Generator().manual_seed(7))
    for holes in (False, True):
        lab = W.make_labels(3, 57, 76, seed=1, holes=holes)
        fx = {"output": out, "labels": lab, "sobel": ref_sobel.Sobel()(out)}
        if not holes:
            fx["loss"] = {k: float(v) for k, v in reference_loss(out, lab).items()}
            mine = rc.sobel_loss_terms(out, lab)
            for k, v in fx["loss"].items():
                assert abs(float(mine[k]) - v) <= 1e-5 * max(1.0, abs(v)), (k, float(mine[k]), v)
        mt = MetricsTracker()
        mt.update(outputs=out, labels=lab)
        fx["metrics"] = {k: float(v) for k, v in mt.to_dict().items()}
        mine = rc.metrics_from_sums(rc.metric_sums(out, lab))
        for k, v in fx["metrics"].items():
            both_nan = mine[k] != mine[k] and v != v          # NaN labels poison mae/mse exactly as in util.py:54-57
            assert both_nan or abs(mine[k] - v) <= 2e-5 * max(1.0, abs(v)), (k, mine[k], v)
        assert torch.allclose(rc.sobel(out), fx["sobel"], atol=1e-5)
        torch.save(fx, os.path.join(HERE, f"sobel_loss_metrics{'_holes' if holes else ''}.pt"))
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
