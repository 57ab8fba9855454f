#!/bin/bash
# multi-GPU bench lines: bash tools/gpu_scale.sh N   (one process per GPU, weak scaling unless noted)
N=${1:-8}
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" --no-cpu-baseline 2> gpurun_out/scale_${tag}_n$N.err | tail -1 > gpurun_out/scale_${tag}_n$N.json; python - "$tag" gpurun_out/scale_${tag}_n$N.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print(sys.argv[1], "n_gpus", d["n_gpus"], "value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], d["config"]["workload"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run enb0hu_b64 --arch enb0-hu --batch 64
run rn50hu_b64 --arch rn50-hu --batch 64
run rn50hu_g256 --arch rn50-hu --batch $((256 / N)) --steps 20
run enb4hu_b32 --arch enb4-hu --batch 32
run enb4lr_b8 --arch enb4-lr --batch 8
