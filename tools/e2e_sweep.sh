for c in 4 8 16 32; do python bench.py --no-cpu-baseline --no-wide --steps 6 --warmup 3 --chunks $c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunks',$c,'value %.4g'%d['value'],'e2e %.4g'%d['e2e']['value'])"; done
