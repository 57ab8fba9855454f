#!/bin/bash
# multi-GPU legs of the second half of round 2 (gpurun --gpus N): the NCCL test of config 4, the default bench line
# (ENB0-HU + RN50-HU, fp32-host and uint8-frame e2e) and config 4 with its all_reduce at N ranks
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu_nccl_gpu.py tests/test_dropin_gpu.py -m gpu -q -s 2>&1 | tail -6 | tee gpurun_out/r02b_pytest_multigpu_n$N.log
run() {  # name, args...
  name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/r02b_scale_${name}_n$N.json 2> gpurun_out/r02b_scale_${name}_n$N.err
  tail -1 gpurun_out/r02b_scale_${name}_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', 'N=$N', d['value'], 'e2e', d['e2e']['value'], d['e2e'].get('u8_frames',{}).get('value'), d.get('loss'), (d.get('rn50_hu') or {}).get('value'), ((d.get('rn50_hu') or {}).get('e2e') or {}).get('value'))"
}
run enb0hu_b64 --steps 20 --warmup 5 --no-cpu-baseline
run cfg4_enb4hu_b32_loss --arch enb4-hu --batch 32 --with-loss --quick
run cfg5_enb4lr_b8 --arch enb4-lr --batch 8 --quick
run cfg3_rn50hu_g256 --arch rn50-hu --batch $((256 / N)) --quick
