python tools/prof_small.py 2>&1 | tail -44
for c in 4 6 8; do python bench.py --no-train --no-small-batch --no-parity --no-cpu-baseline --steps 6 --warmup 3 --e2e-chunks $c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print($c, d['value'], d['e2e']['value'], d['e2e']['ms_per_step'])"; done
