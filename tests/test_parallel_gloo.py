"""not gpu: the N>1 path on CPU -- 2 gloo ranks shard a batch, compute partial sums, all-reduce the
16-double vector, and must reproduce the single-process result exactly (sums are associative in fp64 here)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, json, torch
sys.path.insert(0, os.environ["EDE_ROOT"])
from efficientdepthestimation_b200 import parallel, loss
from oracle import reside_cpu as rc, weights as W
rank, world, _ = parallel.init_from_env("gloo")
out = 0.5 + 9.5 * torch.rand(6, 1, 24, 32, generator=torch.Generator().manual_seed(7))
lab = W.make_labels(6, 24, 32, seed=1)
o, l = parallel.shard_batch(out, rank, world), parallel.shard_batch(lab, rank, world)
# CPU stand-in for the per-rank kernel result: the oracle's sums in the layout of include/ede.h
t, m = rc.sobel_loss_terms(o, l), rc.metric_sums(o, l)
px = float(o.numel())
s = torch.tensor([float(t["depth"]) * px, float(t["normal"]) * px, float(t["dx"]) * px, float(t["dy"]) * px,
                  m["abs"], m["sq"], m["rel"], m["log10"], m["d1"], m["d2"], m["d3"], m["num_valid"], px,
                  float(o.shape[0]), 0.0, 0.0], dtype=torch.float64)
s = parallel.all_reduce_sums(s)
slow = parallel.reduce_max(float(rank + 1))
if rank == 0:
    print(json.dumps({"loss": loss.loss_from_sums(s.tolist()), "metrics": loss.metrics_from_sums(s.tolist()),
                      "world": world, "slow": slow, "bounds": parallel.shard_bounds(6, 1, world)}))
'''


def test_two_rank_scalar_allreduce_matches_single_process(tmp_path):
    import json
    import torch
    from oracle import reside_cpu as rc, weights as W
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, EDE_ROOT=ROOT, OMP_NUM_THREADS="1")
    import socket
    for attempt in range(2):  # a rendezvous port can be taken between probing and binding: retry once
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                            "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                           env=env, capture_output=True, text=True, timeout=300)
        if r.returncode == 0:
            break
    assert r.returncode == 0, r.stderr[-2000:]
    got = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert got["world"] == 2 and got["slow"] == 2.0 and got["bounds"] == [3, 6]
    out = 0.5 + 9.5 * torch.rand(6, 1, 24, 32, generator=torch.Generator().manual_seed(7))
    lab = W.make_labels(6, 24, 32, seed=1)
    ref_loss = {k: float(v) for k, v in rc.sobel_loss_terms(out, lab).items()}
    ref_m = rc.metrics_from_sums(rc.metric_sums(out, lab))
    for k, v in ref_loss.items():
        assert abs(got["loss"][k] - v) <= 1e-5 * max(1.0, abs(v)), k
    for k, v in ref_m.items():
        assert abs(got["metrics"][k] - v) <= 1e-9 * max(1.0, abs(v)), k
