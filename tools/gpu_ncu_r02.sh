#!/bin/bash
# round-2 ncu captures (run on the GPU box through gpurun; outputs under gpurun_out/, summaries are copied to profiles/)
mkdir -p gpurun_out
python bench.py --quick --steps 2 --warmup 3 > gpurun_out/r02_quick.json 2> gpurun_out/r02_quick.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --profile-from-start off -c 800 --csv --log-file gpurun_out/r02_launches.csv python bench.py --quick --steps 2 --warmup 3 > gpurun_out/r02_ncu_launches.log 2>&1
for k in upconv_rows_kernel upconv_cols_kernel tapsum_kernel dwconv_tile_kernel dwconv_tma_kernel se_fc_kernel; do
  timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:$k -c 1 -f -o gpurun_out/r02_prof_$k python bench.py --quick --steps 1 --warmup 3 > gpurun_out/r02_ncu_$k.log 2>&1
done
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -c 1 -f -o gpurun_out/r02_prof_rconv0 python tools/run_op.py enb0-hu 64 R.conv0 1 > gpurun_out/r02_ncu_rconv0.log 2>&1
ls -la gpurun_out/r02_*
