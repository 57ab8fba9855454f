"""not gpu: host logic of the B200 path -- state_dict contract, prepacking, plan compilation (dry, no GPU),
arena layout, error behaviour, sharding arithmetic."""
import io

import pytest
import torch

from efficientdepthestimation_b200 import _cabi as A
from efficientdepthestimation_b200 import engine, factory, loss, parallel
from efficientdepthestimation_b200 import prepack as P
from efficientdepthestimation_b200.engine import compile_module
from oracle import weights as W


@pytest.mark.parametrize("arch", factory.ARCHS)
def test_state_dict_layout_is_the_reference_layout(arch):
    """Keys, order and shapes equal the layout that loads STRICTLY into the real reference modules
    (pinned by tests/golden/make_golden.py)."""
    m = factory.build_model(arch)
    sd = m.state_dict()
    lay = W.layout(arch)
    assert list(sd.keys()) == [k for k, _, _ in lay]
    for k, shape, _ in lay:
        assert tuple(sd[k].shape) == tuple(shape), k


def test_load_state_dict_strict_and_module_prefix_error():
    sd = W.make_state_dict("enb0-hu", seed=1, condition=False)
    m = factory.build_model("enb0-hu")
    m.load_state_dict(sd)                                   # strict by default, as every reference caller does
    bad = {"module." + k: v for k, v in sd.items()}         # DataParallel checkpoints: callers catch RuntimeError
    with pytest.raises(RuntimeError):
        m.load_state_dict(bad)


def test_define_model_and_unsupported_encoders():
    from efficientdepthestimation_b200.ReSIDE.train import define_model
    from efficientdepthestimation_b200.ReSIDE.models import modules
    from efficientdepthestimation_b200.ReSIDE.models.lasinger2019 import Encoder
    m = define_model(is_efficientnet=True, efficientnet_variant="efficientnet-b0")    # no download attempted
    assert m.num_parameters == 5333583
    assert m.E.block_out_channels == [24, 40, 80, 320] and m.E.num_features == 320
    with pytest.raises(RuntimeError, match="Unsupported encoder network"):
        modules.E_efficientnet("vgg16")
    with pytest.raises(RuntimeError, match="Unsupported encoder network"):
        Encoder("densenet161")
    with pytest.raises(RuntimeError):
        define_model(is_densenet=True)


def test_midasnet_checkpoint_roundtrip():
    from efficientdepthestimation_b200.ReSIDE.models.lasinger2019 import MidasNet
    m = factory.build_model("enb0-lr", lr_output_wh=(152, 114))
    assert m.output_size == (152, 114) and m._output_size == (114, 152) and m.input_size == (384, 384)
    f = io.BytesIO()
    m.save(f)
    f.seek(0)
    state = torch.load(f)
    assert set(state) == {"encoder", "decoder", "input_size", "output_size", "adversarial_training", "weights",
                          "version"}
    f.seek(0)
    m2 = MidasNet.load(f)
    assert m2.output_size == (152, 114) and m2.decoder.num_features == 24
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))


def test_forward_has_no_cpu_or_training_fallback():
    m = factory.build_model("enb0-hu")
    with pytest.raises(RuntimeError, match="eval"):
        m(torch.zeros(1, 3, 64, 64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.eval()(torch.zeros(1, 3, 64, 64))
    with pytest.raises(RuntimeError, match="CUDA"):
        loss.loss_metric_sums(torch.zeros(1, 1, 8, 8), torch.ones(1, 1, 8, 8))


@pytest.mark.parametrize("arch,n,expect_gflop", [("enb0-hu", 2, 28.77), ("rn50-hu", 1, 214.05), ("enb4-hu", 1, 40.63)])
def test_dry_plan_compile_counts_the_reference_flops(lib, arch, n, expect_gflop):
    """The plan's algorithmic FLOPs equal the reference graph's 2*MAC -- nothing skipped."""
    m = factory.build_model(arch).eval()
    h, w = factory.input_hw(arch)
    plan, _, _ = compile_module(m, (torch.empty(n, 3, h, w, device="meta"),), torch.device("cpu"), "bf16")
    assert abs(plan.flops / n / 1e9 - expect_gflop) < 0.05 * expect_gflop
    assert plan.outputs[0][1] == (n, 1, 114, 152)
    cfgs = plan.conv_configs()
    assert all(cfg.eligible for _, _, cfg in cfgs), [t for t, _, c in cfgs if not c.eligible]   # all on tcgen05
    assert all(cfg.smem_bytes <= 227 * 1024 and cfg.tmem_cols <= 512 for _, _, cfg in cfgs)
    # every Hu variant takes the low-resolution up-projection path for MFF.up2-4: fewer MACs are issued than the
    # reference graph executes (Graph.upconv)
    assert plan.exec_flops < plan.flops, (arch, plan.exec_flops, plan.flops)
    with pytest.raises(RuntimeError, match="no CPU execution path"):
        plan.run([torch.zeros(n, 3, h, w)])


def test_fp32_plan_uses_exact_path(lib):
    m = factory.build_model("enb0-hu").eval()
    plan, _, _ = compile_module(m, (torch.empty(1, 3, 228, 304, device="meta"),), torch.device("cpu"), "fp32")
    assert not any(cfg.eligible for _, _, cfg in plan.conv_configs())


def test_arena_buffers_never_overlap_while_live():
    m = factory.build_model("enb0-lr").eval()
    g = engine.Graph(torch.device("cpu"), "bf16")
    x = g.input_nchw((2, 3, 384, 384))
    g.output_nchw(engine.emit(m, g, x))
    plan = g.finalize()
    bufs = [b for b in g.bufs if b.first >= 0]
    assert plan.arena_bytes < sum(b.nbytes for b in bufs)          # liveness reuse actually happens
    for i, a in enumerate(bufs):
        assert a.offset % 1024 == 0
        for b in bufs[i + 1:]:
            if a.first <= b.last and b.first <= a.last:            # live at the same time
                assert a.offset + a.nbytes <= b.offset or b.offset + b.nbytes <= a.offset


def test_concat_is_free_and_lane_maps():
    g = engine.Graph(torch.device("cpu"), "bf16")
    d, m = g.alloc_concat(1, 4, 4, [10, 64])
    assert (d.c, d.c_p, d.co, m.co, m.c_p) == (10, 16, 0, 16, 64)
    cat = engine.cat_views([d, m])
    assert cat.c == 74 and cat.c_p == 80 and cat.lanes[:10] == list(range(10)) and cat.lanes[10:16] == [-1] * 6
    assert cat.lanes[16] == 10 and cat.lanes[-1] == 73
    w = torch.arange(74.0).view(1, 74, 1, 1)
    full = P.scatter_weight(w, cat.lanes, [0] + [-1] * 7)
    assert full.shape == (8, 80, 1, 1) and float(full[0, 16, 0, 0]) == 10.0 and float(full[0, 12, 0, 0]) == 0.0


def test_prepack_bn_fold_and_split_weights():
    bn = dict(weight=torch.tensor([2.0]), bias=torch.tensor([0.5]), mean=torch.tensor([1.0]),
              var=torch.tensor([3.0]), eps=1.0)
    sc, sf = P.fold_bn(1, bn, torch.tensor([0.25]), "cpu")
    assert abs(float(sc) - 1.0) < 1e-6 and abs(float(sf) - (0.5 - 1.0 + 0.25)) < 1e-6
    full = torch.randn(8, 8, 3, 3)
    w2, co, ci = P.pack_umma(full, 2)
    assert w2.shape == (18, 16, 64) and (co, ci) == (16, 64)
    rec = (w2[:9].float() + w2[9:].float())[:, :8, :8].reshape(3, 3, 8, 8).permute(2, 3, 0, 1)
    assert float((rec - full).abs().max()) < 2e-5 * float(full.abs().max())     # hi+lo keeps ~16 bits
    assert torch.equal(P.pack_direct(full)[4, :, :], full[:, :, 1, 1].t())


def test_shard_bounds_cover_the_batch():
    for n, world in [(256, 8), (10, 4), (3, 8), (64, 1)]:
        spans = [parallel.shard_bounds(n, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(hi - lo for lo, hi in spans) - min(hi - lo for lo, hi in spans) <= 1


def test_metrics_from_sums_reproduce_reference_quirk():
    s = [0.0] * 16
    s[4], s[5], s[6], s[7], s[8], s[9], s[10], s[11], s[12], s[13] = 20.0, 50.0, 4.0, 3.0, 60, 80, 90, 100.0, 100.0, 4.0
    m = loss.metrics_from_sums(s)
    assert m["mae"] == 0.2 and m["abs_rel"] == 0.04 and m["delta1"] == 0.6
    assert abs(m["log10"] - 3.0 / 100.0 / 4.0) < 1e-12                 # util.py:68 omits the batch factor
    from efficientdepthestimation_b200.ReSIDE.util import MetricsTracker
    t = MetricsTracker()
    t.update_from_sums(s)
    d = t.to_dict()
    assert abs(d["mae"] - 0.2) < 1e-12 and abs(d["log10"] - m["log10"]) < 1e-12 and abs(d["rmse"] - 0.5 ** 0.5) < 1e-12


def test_plan_cache_goes_stale_with_the_weights(lib):
    """Compiled plans hold prepacked COPIES of the weights: any weight change on any submodule must recompile
    (`model.E.load_state_dict`, in-place parameter writes, `set_precision(encoder=...)`)."""
    from efficientdepthestimation_b200 import engine
    m = factory.build_model("enb0-hu").eval()
    x = torch.empty(1, 3, 64, 96)
    p0 = m.plan_for(x)[0]
    assert m.plan_for(x)[0] is p0                                   # cached
    m.E.load_state_dict(m.E.state_dict())                           # submodule load: the PARENT's plan is stale too
    p1 = m.plan_for(x)[0]
    assert p1 is not p0 and m.plan_for(x)[0] is p1
    with torch.no_grad():
        m.R.conv2.bias.add_(1.0)                                    # in-place write on a plain nn.Conv2d child
    p2 = m.plan_for(x)[0]
    assert p2 is not p1
    m.freeze_plans()                                                # opt out of the per-call fingerprint ...
    with torch.no_grad():
        m.R.conv2.bias.add_(1.0)
    assert m.plan_for(x)[0] is p2
    m.D.load_state_dict(m.D.state_dict())                           # ... the epoch still catches load_state_dict
    p3 = m.plan_for(x)[0]
    assert p3 is not p2
    m.freeze_plans(False)
    try:
        engine.set_precision("bf16", encoder="bf16")
        p4 = m.plan_for(x)[0]
        assert p4 is not p3                                         # encoder precision is part of the key
    finally:
        engine.set_precision("bf16", encoder="bf16x2")
    assert m.plan_for(x)[0] is p3


def test_plan_build_packs_on_the_host_and_reserves_device_regions(lib):
    """Dry compile: every conv has one prepack job (host fp32 weight staged, device region reserved), the SE state is
    a zero-initialised region, and nothing of the plan refers to a torch GPU tensor."""
    from efficientdepthestimation_b200.engine import Graph, emit
    m = factory.build_model("enb0-hu").eval()
    g = Graph(torch.device("cpu"), "bf16")
    with torch.no_grad():
        y = emit(m, g, g.input_nchw((1, 3, 228, 304)))
        g.output_nchw(y)
    n_conv = sum(1 for k, *_ in g.ops if k == A.OP_CONV)
    assert len(g.prepacks) == n_conv >= 50
    assert all(j["w_umma"] is not None and j["w_direct"] is None and j["full"].temp for j in g.prepacks)
    assert sum(1 for st in g.stages if st.zero) == 2 * 16            # pool sums + gates of the 16 MBConv blocks
    plan = g.finalize()
    assert plan.weight_bytes > 0 and plan.weights is None and all(st.t is None for st in g.stages)


@pytest.mark.parametrize("name", ["efficientnet-b1", "efficientnet-b3", "efficientnet-b5", "efficientnet-b6",
                                  "efficientnet-b7"])
def test_every_efficientnet_split_compiles_to_a_plan(lib, name):
    """The B1..B7 block splits of ReSIDE/models/modules.py:168-181 (B0 / B4 are covered by the parity tests): the
    encoder builds with the reference's channel counts and the whole Hu model compiles onto the tcgen05 path."""
    from oracle import effnet_spec as spec
    from efficientdepthestimation_b200.ReSIDE.models import modules, net
    enc = modules.E_efficientnet(name, pretrained=False)
    _, blocks, _ = spec.block_specs(name)
    idx = spec.split_indices(name)
    assert enc.block_out_channels == [blocks[i - 1].cout for i in idx[1:]]
    assert [len(b) for b in enc.blocks] == [idx[1] + 2] + [b - a for a, b in zip(idx[1:-1], idx[2:])]
    m = net.model(enc, enc.num_features, enc.block_out_channels).eval()
    plan, _, _ = compile_module(m, (torch.empty(1, 3, 228, 304, device="meta"),), torch.device("cpu"), "bf16")
    assert plan.outputs[0][1] == (1, 1, 114, 152)
    cfgs = plan.conv_configs()
    assert all(cfg.eligible for _, _, cfg in cfgs), [t for t, _, c in cfgs if not c.eligible]
    assert all(cfg.smem_bytes <= 227 * 1024 and cfg.tmem_cols <= 512 for _, _, cfg in cfgs)


def test_unknown_efficientnet_name_raises_like_the_reference():
    from efficientdepthestimation_b200.ReSIDE.models import modules
    with pytest.raises(RuntimeError, match="Unsupported encoder network"):
        modules.E_efficientnet("efficientnet-b9", pretrained=False)
    with pytest.raises(RuntimeError, match="Unsupported encoder network"):
        modules.E_efficientnet("resnet50", pretrained=False)


def test_bottleneck_decoder_layout_and_plan(lib):
    """Decoder(block_type=BottleneckBlock): state_dict keys equal the layout pinned against the reference
    (tests/golden/make_golden_bottleneck.py), strict load works, the plan compiles."""
    from oracle import weights as W
    m = factory.build_model("enb0-lr", lr_block="bottleneck")
    sd = W.make_state_dict("enb0-lr", seed=4321, lr_block="bottleneck", condition=False)
    assert list(m.state_dict().keys()) == list(sd.keys())
    assert all(tuple(v.shape) == tuple(sd[k].shape) for k, v in m.state_dict().items())
    m.load_state_dict(sd, strict=True)
    plan, _, _ = compile_module(m.eval(), (torch.empty(1, 3, 384, 384, device="meta"),), torch.device("cpu"), "bf16")
    assert plan.outputs[0][1] == (1, 1, 192, 192)
    assert sum(1 for t in plan.tags if t.endswith(".conv3")) >= 7        # 4 res_block1 + 3 res_block2 bottlenecks
