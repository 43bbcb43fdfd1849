cd $GRAFT_REPO_ROOT
for cfg in "--scene disc --tree bvhref --n-per-gpu 1000000" "--scene disc --tree lbvh --n-per-gpu 1000000" "--scene plummer --tree bvhref --n-per-gpu 4000000" "--scene default --tree bvhref --theta 50" "--scene default --tree bvhref --theta 0.5"; do
  echo "== $cfg"
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e $cfg 2>&1 | tail -1 | python -c "
import sys, json
try:
    d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','phases_ms','interactions_per_step','gpu_launches')}, d['roofline']['frac'])
except Exception as e: print('ERR', e)"
done
