# gpurun helper: the end-of-round-2 record on one GPU (tests, ncu launch list with DRAM bytes, both bench arms)
cd $GRAFT_REPO_ROOT
TAG=${1:-r02i}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/tests_$TAG.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tests_$TAG.log
bash scripts/r02_ncu_kernels.sh ${TAG}_1m
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/ref_$TAG.json 2> gpurun_out/ref_$TAG.err; echo ref rc=$?
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo bench rc=$?
NB_B200_NO_REPAIR=1 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_${TAG}_norepair.json 2> gpurun_out/bench_${TAG}_norepair.err; echo bench-norepair rc=$?
python bench.py > gpurun_out/bench_${TAG}_default.json 2> gpurun_out/bench_${TAG}_default.err; echo bench-default rc=$?
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/bench_r02i*.json')) + ['gpurun_out/ref_r02i.json']:
    try:
        d = [json.loads(l) for l in open(f) if l.startswith('{')][-1]
        print(f, d.get('steps'), round(d['ms_per_step'], 4), d.get('phases_ms'), '%.3e' % d['value'], (d.get('roofline') or {}).get('frac'), (d.get('e2e') or {}).get('ms_per_step'), d.get('gpu_launches'))
    except Exception as e:
        print(f, 'ERR', e)
PY
