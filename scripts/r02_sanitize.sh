# compute-sanitizer on one small step of every tree kind (ONE tool per gpurun call, see B200_PROFILING.md):
#   gpurun -- 'bash scripts/r02_sanitize.sh memcheck'      gpurun -- 'bash scripts/r02_sanitize.sh racecheck'
cd $GRAFT_REPO_ROOT
TOOL=${1:-memcheck}
python scripts/sanitize_case.py 100000 > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python scripts/sanitize_case.py 100000 > gpurun_out/r02_sanitize_$TOOL.log 2>&1
echo "rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize case done" gpurun_out/r02_sanitize_$TOOL.log; tail -3 gpurun_out/r02_sanitize_$TOOL.log
