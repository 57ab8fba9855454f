#!/bin/bash
timeout 600 python bench.py --arch rn50-hu --batch 64 --quick --steps 40 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['clocks'])"
timeout 600 python bench.py --quick --steps 40 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['clocks'])"
