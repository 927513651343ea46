#!/bin/bash
# One multi-GPU box session: device-1 / pool tests, the box's H2D ceiling, the single-process pool sweep and torchrun bench runs.
# usage: tools/multi_gpu_session.sh <max gpus> "<config:N config:N ...>"   outputs under gpurun_out/r02_mg_*
G=$1; RUNS=$2; OUT=gpurun_out
python -m pytest tests/test_multi_device.py tests/test_pool.py tests/test_shard.py -m gpu -q 2>&1 | tail -4 > $OUT/r02_mg_tests.log
./tools/h2d_ceiling 20 0.5 > $OUT/r02_h2d_ceiling_${G}gpu.json 2>> $OUT/r02_mg_tests.log
python tools/pool_scale.py --config 0 --steps 4 > $OUT/r02_pool_scale_cfg0.json 2> $OUT/r02_pool_scale_cfg0.err
PORT=29610
for r in $RUNS; do
  c=${r%%:*}; n=${r##*:}; PORT=$((PORT+1))
  if [ "$n" = "1" ]; then
    python bench.py --config $c --steps 10 --no-cpu > $OUT/r02_scale_cfg${c}_n1.json 2> $OUT/r02_scale_cfg${c}_n1.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n --config $c --steps 10 --no-cpu \
      > $OUT/r02_scale_cfg${c}_n$n.json 2> $OUT/r02_scale_cfg${c}_n$n.err
  fi
  echo "cfg $c n $n rc=$?" >> $OUT/r02_mg_tests.log
done
