#!/bin/bash
# round-2 (second half) validation + the measurements committed under profiles/ (one GPU)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -2 | tee gpurun_out/r02b_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r02b_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/r02b_smoke.log
timeout 900 python bench.py > gpurun_out/r02b_bench_enb0hu_b64.json 2> gpurun_out/r02b_bench.err; cut -c1-300 gpurun_out/r02b_bench_enb0hu_b64.json
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r02b_bench_reference_arm.json 2>> gpurun_out/r02b_bench.err
timeout 600 python bench.py --arch enb4-hu --batch 32 --with-loss --quick > gpurun_out/r02b_bench_cfg4_n1.json 2>> gpurun_out/r02b_bench.err
timeout 600 python bench.py --arch enb4-lr --batch 8 --quick > gpurun_out/r02b_bench_cfg5_n1.json 2>> gpurun_out/r02b_bench.err
timeout 600 python bench.py --arch rn50-hu --batch 256 --quick > gpurun_out/r02b_bench_cfg3_n1.json 2>> gpurun_out/r02b_bench.err
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/r02b_plan_enb0hu_b64.txt 2>&1
timeout 300 python tools/profile_plan.py rn50-hu 64 --top 100 > gpurun_out/r02b_plan_rn50hu_b64.txt 2>&1
tail -3 gpurun_out/r02b_bench.err
# ncu: launch list with DRAM bytes (same command as the quick bench), one full capture of the new kernel
python bench.py --quick --steps 2 --warmup 3 > gpurun_out/r02b_quick.json 2> gpurun_out/r02b_quick.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --profile-from-start off -c 800 --csv --log-file gpurun_out/r02b_launches.csv python bench.py --quick --steps 2 --warmup 3 > gpurun_out/r02b_ncu_launches.log 2>&1
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:conv_npack -c 1 -f -o gpurun_out/r02b_prof_npack_dup4 python tools/run_op.py enb0-hu 64 D.up4.conv1+conv2 1 > gpurun_out/r02b_ncu_npack.log 2>&1
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:conv_npack -c 1 -f -o gpurun_out/r02b_prof_npack_c12 python tools/run_op.py enb0-hu 64 MFF.up2.conv1_2 1 >> gpurun_out/r02b_ncu_npack.log 2>&1
ls -la gpurun_out/r02b_* | tail -20
