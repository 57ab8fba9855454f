#!/bin/bash
# multi-GPU legs (gpurun --gpus N): NCCL test of config 4, then the scaling lines that go to profiles/r02_scaling.jsonl
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu_nccl_gpu.py tests/test_dropin_gpu.py -m gpu -q -s 2>&1 | tail -12 | tee gpurun_out/r02_pytest_multigpu_n$N.log
run() {  # name, nproc, args...
  name=$1; np=$2; shift 2
  if [ "$np" = "1" ]; then timeout 900 python bench.py --gpus 1 "$@" > gpurun_out/r02_scale_${name}_n1.json 2> gpurun_out/r02_scale_${name}_n1.err
  else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $np "$@" > gpurun_out/r02_scale_${name}_n$np.json 2> gpurun_out/r02_scale_${name}_n$np.err; fi
  tail -1 gpurun_out/r02_scale_${name}_n$np.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', 'N=$np', d['value'], 'e2e', d['e2e']['value'], d['e2e'].get('u8_frames',{}).get('value'), d.get('loss'), (d.get('rn50_hu') or {}).get('value'))"
}
for np in 1 $N; do
  run enb0hu_b64 $np --steps 20 --warmup 5 --no-cpu-baseline
  run cfg4_enb4hu_b32_loss $np --arch enb4-hu --batch 32 --with-loss --quick
  run cfg5_enb4lr_b8 $np --arch enb4-lr --batch 8 --quick
done
run cfg3_rn50hu_g256 $N --arch rn50-hu --batch $((256 / N)) --quick
