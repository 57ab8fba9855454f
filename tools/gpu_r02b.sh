#!/bin/bash
# round-2 (second half): the no-alloc test alone in a fresh process, then the whole GPU suite, smoke, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -3 | tee gpurun_out/r02b_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r02b_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/r02b_smoke.log
timeout 900 python bench.py > gpurun_out/r02b_bench_enb0hu_b64.json 2> gpurun_out/r02b_bench.err; cut -c1-400 gpurun_out/r02b_bench_enb0hu_b64.json
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/r02b_plan_enb0hu_b64.txt 2>&1
tail -3 gpurun_out/r02b_bench.err
