#!/bin/bash
# usage (on the GPU box): tools/prof_tractor.sh <tag>
# plain run first (must exit 0), then one full ncu capture of k_tractor with source counters: launch 48 = the 17th of
# the 32 frames bench.py renders at N=1 (timeline position 128, the middle of its window)
T=${1:-rX}; O=gpurun_out; mkdir -p $O
python tools/ktractor.py --steps 10 > $O/${T}_ktractor.txt 2>&1 || { cat $O/${T}_ktractor.txt; exit 1; }
cat $O/${T}_ktractor.txt
python tools/ktractor.py --steps 1 --distinct 32 --repeat 1 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_tractor -s 48 -c 1 -o $O/${T}_k_tractor -f \
    python tools/ktractor.py --steps 1 --distinct 32 --repeat 1 > $O/${T}_ncu.log 2>&1
tail -3 $O/${T}_ncu.log
ncu -i $O/${T}_k_tractor.ncu-rep --page details > $O/${T}_k_tractor_details.txt 2>&1
python tools/ncusum.py $O/${T}_k_tractor.ncu-rep > $O/${T}_k_tractor_summary.txt 2>&1
cat $O/${T}_k_tractor_summary.txt
