#!/bin/bash
# DRAM bytes of the fused kernels for a few zero-skipping policies.
O=gpurun_out/${1:-nd}; mkdir -p $O
for z in 1 1000; do
  for t in 64 32; do
    ASSET_B200_FUSED_ZRUN=$z ASSET_B200_FUSED_THREADS=$t timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_write.sum --clock-control none -k regex:ab_fused -c 2 --csv --log-file $O/z${z}_t${t}.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e > /dev/null 2>&1
    echo "zrun=$z threads=$t"; grep -E "ab_fused" $O/z${z}_t${t}.csv | awk -F'","' '{print $5, $(NF-2), $(NF-1), $NF}' | tr -d '"'
  done
done
