#!/bin/bash
# round-2 validation + the measurements that are committed under profiles/ (one GPU)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q 2>&1 | tail -5 | tee gpurun_out/r02_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/r02_smoke.log
timeout 900 python bench.py > gpurun_out/r02_bench_enb0hu_b64.json 2> gpurun_out/r02_bench.err; cut -c1-300 gpurun_out/r02_bench_enb0hu_b64.json
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r02_bench_reference_arm.json 2>> gpurun_out/r02_bench.err
timeout 600 python bench.py --arch enb4-hu --batch 32 --with-loss --quick > gpurun_out/r02_bench_cfg4_n1.json 2>> gpurun_out/r02_bench.err
timeout 600 python bench.py --arch enb4-lr --batch 8 --quick > gpurun_out/r02_bench_cfg5_n1.json 2>> gpurun_out/r02_bench.err
timeout 600 python bench.py --arch rn50-hu --batch 256 --quick > gpurun_out/r02_bench_cfg3_n1.json 2>> gpurun_out/r02_bench.err
timeout 300 python tools/bench_render.py > gpurun_out/r02_render.json 2>> gpurun_out/r02_bench.err; cat gpurun_out/r02_render.json
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/r02_plan_enb0hu_b64.txt 2>&1
timeout 300 python tools/profile_plan.py rn50-hu 64 --top 100 > gpurun_out/r02_plan_rn50hu_b64.txt 2>&1
timeout 300 python tools/bench_pipeline.py > gpurun_out/r02_pipeline.jsonl 2>> gpurun_out/r02_bench.err
tail -3 gpurun_out/r02_bench.err
