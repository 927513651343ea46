#!/bin/bash
# A/B kernel builds on the GPU box: tools/ab.sh lib1.so lib2.so ... (paths relative to orb_slam2v2-1_b200/orbx)
for so in "$@"; do
  ORBX_LIB=$PWD/orb_slam2v2-1_b200/orbx/$so python bench.py --no-cpu --steps 10 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$so', round(d['value']), round(d['e2e']['value']), {k: round(v,3) for k,v in d['roofline']['stage_ms'].items()})"
done
