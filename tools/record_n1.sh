#!/bin/bash
# bench lines of the four configurations at N = 1 (with the CPU baseline) and the reference arm: gpurun_out/r02_scale_cfg<c>_n1.json
for c in ${1:-0 1 2 4}; do
  python bench.py --config $c > gpurun_out/r02_scale_cfg${c}_n1.json 2> gpurun_out/r02_scale_cfg${c}_n1.err; echo "cfg $c rc=$?"
done
python bench.py --impl reference --config 0 --steps 3 --warmup 1 > gpurun_out/r02_reference_cfg0.json 2> gpurun_out/r02_reference_cfg0.err; echo "ref rc=$?"
