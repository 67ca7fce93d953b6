#!/bin/bash
# usage (on the GPU box): tools/profile_weak.sh <tag>
# One `ncu --set full` capture per kbench row that is still well under the HBM roof (the round-2 list in
# DESIGN.md §3.1), each after the same kbench command has exited 0 without ncu.  Writes the details page
# and tools/ncusum.py's one-screen summary under gpurun_out/<tag>_weak_<n>.*
T=${1:-rX}; O=gpurun_out; mkdir -p $O
i=0
while IFS='|' read -r name kern; do
  i=$((i+1))
  python tools/kbench.py --only "$name" --reps 2 > $O/${T}_weak_$i.kbench.txt 2>&1 || continue
  ncu --set full --clock-control none --import-source on -k "regex:$kern" -s 3 -c 1 -o $O/${T}_weak_$i -f \
      python tools/kbench.py --only "$name" --reps 2 > /dev/null 2>&1
  { echo "# kbench row: $name"; tail -1 $O/${T}_weak_$i.kbench.txt; python tools/ncusum.py $O/${T}_weak_$i.ncu-rep; } > $O/${T}_weak_$i.summary.txt 2>&1
  ncu -i $O/${T}_weak_$i.ncu-rep --page details > $O/${T}_weak_$i.details.txt 2>&1
done <<'ROWS'
K2 composite 50% sub-rect at odd x|k_composite_rows
K2 composite 70% sub-rect odd x +alpha_a+alpha_b|k_composite_rows
K2 composite luma wipe (scalar)|k_composite_rows
K3 luma wipe yuv422 translucent|k_luma_yuv422
K4 affine rgba bilinear|k_affine
K4 affine rgba bicubic|k_affine
K4 affine rgba nearest|k_affine
ROWS
cat $O/${T}_weak_*.summary.txt | head -150
