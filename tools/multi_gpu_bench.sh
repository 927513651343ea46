#!/bin/bash
# torchrun bench runs only (no tests, no ceilings): tools/multi_gpu_bench.sh "<config:N config:N ...>"   outputs gpurun_out/r02_scale_cfg<c>_n<N>.json
OUT=gpurun_out; PORT=29710
for r in $1; do
  c=${r%%:*}; n=${r##*:}; PORT=$((PORT+1))
  if [ "$n" = "1" ]; then
    python bench.py --config $c --steps 10 --no-cpu > $OUT/r02_scale_cfg${c}_n1.json 2> $OUT/r02_scale_cfg${c}_n1.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n --config $c --steps 10 --no-cpu \
      > $OUT/r02_scale_cfg${c}_n$n.json 2> $OUT/r02_scale_cfg${c}_n$n.err
  fi
  echo "cfg $c n $n rc=$? $(tail -c 300 $OUT/r02_scale_cfg${c}_n$n.json | head -c 0)"
done
