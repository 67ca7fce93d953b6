#!/bin/bash
# usage (on the GPU box): tools/profile_round.sh <tag>   e.g. r2b
# Everything profiles/ holds for one round: GPU tests, both bench arms, per-kernel table, launch list of
# the bench command's device legs, one full ncu capture of the fused kernel (tools/prof_tractor.sh), the melt-level
# consumer comparison.  Each profiler run comes after the same command has exited 0 without ncu.
T=${1:-rX}; O=gpurun_out
mkdir -p $O
(nvidia-smi -L; nproc; lscpu | grep "Model name") > $O/${T}_box.txt 2>&1
python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/${T}_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.txt 2>&1
python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err || exit 1
python bench.py --impl reference --steps 10 --warmup 3 > $O/${T}_bench_reference.json 2>> $O/${T}_bench.err
python tools/kbench.py --out $O/${T}_kbench.jsonl > /dev/null 2>&1
python bench.py --steps 2 --warmup 3 --repeat 1 --no-graph --no-melt --no-cpu-baseline --no-e2e-abi --no-extra > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --repeat 1 --no-graph --no-melt --no-cpu-baseline --no-e2e-abi --no-extra > /dev/null 2>&1
tools/prof_tractor.sh $T > $O/${T}_prof_tractor.txt 2>&1
python tools/melt_consumers.py 8 12000 > $O/${T}_melt_consumers.txt 2>&1
tail -2 $O/${T}_pytest.txt; cat $O/${T}_smoke.txt; cat $O/${T}_bench.json; tail -4 $O/${T}_melt_consumers.txt
