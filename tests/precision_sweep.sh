#!/bin/bash
# error (tests/precision_probe.py) and step time of the accurate-D levels (diagnostic, GPU box)
for lv in 0 1 2 3; do
  echo "== EDE_ACCURATE_D=$lv"
  EDE_ACCURATE_D=$lv python tests/precision_probe.py enb4-hu randn 32 | grep "bf16/bf16x2"
  EDE_ACCURATE_D=$lv python tests/precision_probe.py enb0-hu randn 64 | grep "bf16/bf16x2"
  EDE_ACCURATE_D=$lv python bench.py --steps 10 --warmup 3 --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('enb0-hu b64', d['value'], 'maps/s', d['ms_per_step'], 'ms')"
  EDE_ACCURATE_D=$lv python bench.py --arch enb4-hu --batch 32 --steps 10 --warmup 3 --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('enb4-hu b32', d['value'], 'maps/s', d['ms_per_step'], 'ms')"
done
