#!/bin/bash
# The evidence of a round on a 1-GPU box, in the order the profiling recipe asks for: every program runs once WITHOUT ncu before any
# capture.  usage: tools/record_round.sh TAG     outputs: gpurun_out/TAG_*   (then, here: python tools/make_profiles.py TAG "note")
T=${1:-r02}; O=gpurun_out
python tools/profile_once.py 512 > $O/${T}_profile_once.log 2>&1 || exit 1
ncu --set full --import-source on --clock-control none --launch-skip 22 --launch-count 22 -o $O/${T}_full -f python tools/profile_once.py 512 > $O/${T}_ncu_full.log 2>&1
python bench.py --steps 2 --warmup 1 --no-cpu > $O/${T}_b.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches_b512.csv python bench.py --steps 2 --warmup 1 --no-cpu > $O/${T}_ncu_launch.log 2>&1
python tools/bench_latency.py > $O/${T}_latency.jsonl 2> $O/${T}_latency.err
python tools/bench_map_ops.py > $O/${T}_map_ops.jsonl 2> $O/${T}_map_ops.err
tail -2 $O/${T}_ncu_full.log; wc -l $O/${T}_launches_b512.csv $O/${T}_latency.jsonl $O/${T}_map_ops.jsonl
