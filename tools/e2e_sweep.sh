for b in 256 512 1024 2048; do
  python bench.py --no-cpu --steps 6 --batch $b 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('batch $b e2e', round(d['e2e']['value']), 'frac', round(d['e2e']['ceiling']['frac'],3), 'value', round(d['value']), d['e2e']['how'][:90])"
done
