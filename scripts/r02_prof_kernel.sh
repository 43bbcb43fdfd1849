# gpurun helper: one full ncu capture of a named kernel inside a step at N bodies
#   gpurun -- 'bash scripts/r02_prof_kernel.sh sort_onesweep_kernel 8000000 tag [skip]'
cd "${GRAFT_REPO_ROOT:-.}"
K=$1; N=${2:-1000000}; TAG=${3:-r02k}; SKIP=${4:-12}
python scripts/ab_step.py --n $N --reps 1 --steps 3 > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -f \
    -o gpurun_out/prof_${TAG}_$K python scripts/ab_step.py --n $N --reps 1 --steps 3 > gpurun_out/ncu_$TAG.log 2>&1
echo "rc=$?"
