#!/usr/bin/env bash
# one gpurun call: plain run, launch list of a whole eager step, ncu --set full of the conv / pooling kernels of one step
mkdir -p gpurun_out
python scripts/ncu_target.py > gpurun_out/target_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 326 -c 170 --csv --log-file gpurun_out/r02_launches_b.csv python scripts/ncu_target.py > gpurun_out/ncu_list.log 2>&1
python scripts/ncu_target.py > gpurun_out/target_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tc_conv2_kernel|tc_pool_kernel" -s 48 -c 24 -o gpurun_out/r02_conv_pool -f python scripts/ncu_target.py > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/r02_conv_pool.ncu-rep --page raw --csv > gpurun_out/r02_conv_pool_raw.csv 2>/dev/null
ls -la gpurun_out/r02_conv_pool.ncu-rep
if [ $(stat -c %s gpurun_out/r02_conv_pool.ncu-rep) -gt 45000000 ]; then rm gpurun_out/r02_conv_pool.ncu-rep; fi
tail -2 gpurun_out/ncu_full.log
