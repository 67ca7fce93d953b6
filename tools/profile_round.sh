#!/bin/bash
# usage (on the GPU box): tools/profile_round.sh <tag>   e.g. r1k
# Everything profiles/ holds for one round: GPU tests, both bench arms, per-kernel table, launch list of
# the bench command, one full ncu capture of the fused kernel, melt-level comparison.  Each profiler run
# comes after the same command has exited 0 without ncu.
T=${1:-rX}; O=gpurun_out
mkdir -p $O
(nvidia-smi -L; nproc; lscpu | grep "Model name") > $O/${T}_box.txt 2>&1
python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/${T}_pytest.txt
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_reference.json 2>> $O/${T}_bench.err
python tools/kbench.py --out $O/${T}_kbench.jsonl > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 2 --warmup 1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tractor -s 40 -c 2 -o $O/${T}_k_tractor -f \
    python bench.py --steps 2 --warmup 1 > /dev/null 2>&1
ncu -i $O/${T}_k_tractor.ncu-rep --page details > $O/${T}_k_tractor_ncu_full_details.txt 2>&1
python tools/melt_bench.py > $O/${T}_melt_bench.txt 2>&1
tail -2 $O/${T}_pytest.txt; cat $O/${T}_bench.json; tail -2 $O/${T}_melt_bench.txt
