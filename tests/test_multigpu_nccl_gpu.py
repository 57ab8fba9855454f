"""-m gpu: BASELINE.json config 4 AS WRITTEN, on real devices: ENB4-HU forward -> fused resize + Sobel loss + metric
sums (ede_depth_epilogue, kernel-produced fp64 sums on the DEVICE) -> ONE all_reduce(SUM) of the 16-double vector
over NCCL -> rank 0 turns the global sums into loss / REL / DEL1 (ReSIDE/train.py:216-236,266-278, util.py:39-83).

One process per GPU under torchrun; the batch is sharded on dim 0, no collective on the forward path.  Checked on
rank 0 against the single-process result over the concatenated batch: counters exact, sums <= 1e-12 relative.
world_size 2 needs two GPUs (skipped otherwise; run with `gpurun --gpus 2`); world_size 1 runs on any GPU box and
covers the same code path without the collective.
"""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.environ["EDE_ROOT"])
from efficientdepthestimation_b200 import factory, loss, parallel
from oracle import weights as W            # test infrastructure: seeded weights / inputs / labels only

rank, world, local = parallel.init_from_env()
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
arch, per = os.environ.get("EDE_ARCH", "enb4-hu"), int(os.environ.get("EDE_PER_RANK", "4"))
sd = W.make_state_dict(arch, seed=1234, calib_batch=1)
x = W.make_input(arch, per * world, seed=0)
lab = W.make_labels(per * world, 228, 304, seed=1)
model = factory.build_model(arch)
model.load_state_dict(sd, strict=True)
model = model.to(dev).eval()
xs, ls = parallel.shard_batch(x, rank, world).to(dev), parallel.shard_batch(lab, rank, world).to(dev)
with torch.no_grad():
    y = model(xs)                                                     # no collective on the forward path
    ep = loss.depth_epilogue(y, label=ls, want_sums=True, want_edges=True)
sums = parallel.all_reduce_sums(ep["sums"])                           # DEVICE fp64[16] over NCCL
assert sums.is_cuda and sums.dtype == torch.float64
edges = parallel.all_reduce_sums(ep["edges"])
if world > 1:                                                         # test-only: collect the predictions on rank 0
    ys = [torch.empty_like(y) for _ in range(world)]
    dist.all_gather(ys, y)
    y_all = torch.cat(ys)
else:
    y_all = y
if rank == 0:
    with torch.no_grad():
        single = loss.depth_epilogue(y_all, label=lab.to(dev), want_sums=True, want_edges=True)
    torch.cuda.synchronize()
    s, s1 = sums.cpu().tolist(), single["sums"].cpu().tolist()
    print(json.dumps({"world": world, "backend": dist.get_backend() if world > 1 else "none", "sums": s, "single": s1,
                      "edges": edges.cpu().tolist(), "edges_single": single["edges"].cpu().tolist(),
                      "loss": loss.loss_from_sums(s), "metrics": loss.metrics_from_sums(s),
                      "devices": torch.cuda.device_count()}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
'''


@pytest.mark.parametrize("world", [1, 2])
def test_config4_forward_loss_allreduce(cuda, lib, tmp_path, world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {torch.cuda.device_count()}")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, EDE_ROOT=ROOT, OMP_NUM_THREADS="4")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    got = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert got["world"] == world and (world == 1 or got["backend"] == "nccl")
    s, s1 = got["sums"], got["single"]
    for k in (8, 9, 10, 11, 12, 13):                       # counters: #delta<1.25^k, #valid, #pixels, #images
        assert s[k] == s1[k], (k, s[k], s1[k])
    assert s[12] == 4 * world * 228 * 304 and s[13] == 4 * world
    for k in range(8):                                     # fp64 sums of fp32 terms
        if s1[k] != s1[k]:                                 # log10 of a non-positive prediction: NaN, as in util.py:65-68
            assert s[k] != s[k], (k, s[k], s1[k])
            continue
        assert abs(s[k] - s1[k]) <= 1e-12 * max(1.0, abs(s1[k])), (k, s[k], s1[k])
    assert got["edges"] == got["edges_single"]
    assert got["loss"]["loss"] == got["loss"]["loss"] and 0.0 < got["metrics"]["delta1"] <= 1.0
    print(f"\nconfig 4, {world} rank(s): loss {got['loss']['loss']:.6f} REL {got['metrics']['abs_rel']:.4f} "
          f"DEL1 {got['metrics']['delta1']:.4f}")
