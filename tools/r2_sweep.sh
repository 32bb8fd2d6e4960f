#!/bin/bash
# mid-flow (48x72x728) and 96x144x728 streaming-kernel knob sweep; one process per setting
for shape in 3 2; do
for cap in 16 8 4 2; do for rows in 4 8 16; do
  echo "== shape $shape BGR_EW_CAP=$cap BGR_EW_ROWS=$rows"
  BGR_EW_CAP=$cap BGR_EW_ROWS=$rows python tools/dev_kernel_bench.py --only $shape 2>&1 | grep -E "bn_apply|bn_backward"
done; done
for bps in 2 3; do
  echo "== shape $shape BGR_RED_BPS_WIDE=$bps"
  BGR_RED_BPS_WIDE=$bps python tools/dev_kernel_bench.py --only $shape 2>&1 | grep -E "bn_stats|bn_backward"
done
done
