#!/bin/bash
# full GPU validation + the measurements that are committed under profiles/
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench_enb0.log 2>&1; tail -1 gpurun_out/bench_enb0.log
timeout 600 python bench.py --arch rn50-hu --no-cpu-baseline > gpurun_out/bench_rn50.log 2>&1; tail -1 gpurun_out/bench_rn50.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --profile-from-start off -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1; tail -1 gpurun_out/ncu1.log | cut -c1-200
timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -c 1 -f -o gpurun_out/prof_rconv0 python tools/run_op.py enb0-hu 64 R.conv0 1 > gpurun_out/ncu_rconv0.log 2>&1; tail -1 gpurun_out/ncu_rconv0.log
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/prof_enb0_full.log 2>&1
timeout 300 python tools/profile_plan.py rn50-hu 32 --top 100 > gpurun_out/prof_rn50_full.log 2>&1
timeout 120 tools/micro/mma_contend > gpurun_out/mma_contend.txt 2>&1
timeout 600 python tools/bench_pipeline.py > gpurun_out/pipeline.jsonl 2> gpurun_out/pipeline.err; cat gpurun_out/pipeline.jsonl | cut -c1-220
for k in preprocess_kernel depth_epilogue_kernel depth_resize_convert_kernel; do timeout 300 ncu --set full --import-source on --clock-control none -k regex:$k -s 5 -c 1 -f -o gpurun_out/prof_$k python tools/bench_pipeline.py --no-cpu --kernels-only --steps 4 --warmup 4 > gpurun_out/ncu_$k.log 2>&1; done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:stem_conv -s 3 -c 1 -f -o gpurun_out/prof_stem_rn50 python bench.py --arch rn50-hu --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_stem_rn50.log 2>&1
timeout 300 python tools/profile_plan.py rn50-hu 64 --top 100 > gpurun_out/prof_rn50_b64.log 2>&1
