#!/bin/sh
# One box session: GPU test suite, both bench arms, the ncu launch list of the bench command and the ncu --set full capture of
# the chain kernel (DRAM traffic per launch).  Outputs under gpurun_out/ with the given tag.  Usage: scripts/profile_round.sh r02_v2
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
python -m pytest tests -x -q -m gpu > $O/${TAG}_gputest.log 2>&1; tail -3 $O/${TAG}_gputest.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err || tail -5 $O/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_ref.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --skip-extras > $O/${TAG}_ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:chain2_kernel -s 9 -c 3 -o $O/${TAG}_chain2 python bench.py --steps 2 --warmup 3 --skip-extras > $O/${TAG}_ncu_full.log 2>&1
ncu -i $O/${TAG}_chain2.ncu-rep --page raw --csv > $O/${TAG}_chain2_raw.csv 2>/dev/null
ls -la $O | tail -12
