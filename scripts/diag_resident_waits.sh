# Wait-time profile of the persistent loss+gradient launch: instrumented build, run, normal build.
cd "$(dirname "$0")/.."
(cd hand-integral-pose-estimation_b200 && HIPE_NVCC_EXTRA="-DHIPE_RES_PROF $HIPE_PROF_EXTRA" python -m hipe_b200._build --force) > /dev/null 2>&1
HIPE_NO_BUILD=1 python scripts/diag_resident_waits.py "$@"
(cd hand-integral-pose-estimation_b200 && python -m hipe_b200._build --force) > /dev/null 2>&1
