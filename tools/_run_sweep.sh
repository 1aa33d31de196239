run() { python bench.py --no-cpu-baseline --steps 10 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('%.1f Mrays/s e2e %.1f fallback %d'%(d['value'], d['e2e']['value'], d['fallback_segments']))"; }
echo "sweep all, weight tris"; run
echo "weight 1"; OKRT_TOP_WEIGHT=1 run
echo "weight tris+2"; OKRT_TOP_WEIGHT=2 run
echo "sweep off"; OKRT_TOP_SWEEP=0 run
