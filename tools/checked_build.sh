#!/bin/bash
# The repo's own memcheck (compute-sanitizer is closed on this GPU pool):   bash tools/checked_build.sh [tag]
# Builds liborbfe.so with -DORBFE_CHECKS (device-side assertions on every data-derived shared-memory index / global
# offset of the extraction and stereo kernels, 4 KiB guard bands around the device arenas, verified when a handle is
# destroyed), runs the whole GPU test suite and smoke() against it, then restores the normal build.  A failed
# assertion prints the expression and traps (the CUDA error fails the test); a damaged guard band aborts the process.
# Run it on the GPU box (under gpurun); the log lands in gpurun_out/checked_<tag>.log.
set -u
TAG=${1:-r2}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/gpurun_out
mkdir -p $OUT
cd $ROOT
cp my_orb_slam2_b200/liborbfe.so /tmp/liborbfe_normal.so
if [ -f tools/scratch/var/checked.so ]; then
  cp tools/scratch/var/checked.so my_orb_slam2_b200/liborbfe.so          # prebuilt where nvcc is (the snapshot carries it)
else
  touch my_orb_slam2_b200/csrc/*.cu && make -C my_orb_slam2_b200/csrc -j8 EXTRA=-DORBFE_CHECKS > $OUT/checked_build_$TAG.log 2>&1 || { echo "checked build failed"; exit 1; }
fi
{
  echo "== checked build (-DORBFE_CHECKS): pytest -m gpu"
  python -m pytest tests -m gpu -x -q 2>&1 | tail -5
  echo "== checked build: randomised sweeps (tests/fuzz_parity.py, tests/fuzz_matchers.py)"
  python tests/fuzz_parity.py --first ${FUZZ_FIRST:-5000} --count ${FUZZ_COUNT:-300} 2>&1 | grep -v "refused:" | tail -3
  python tests/fuzz_matchers.py --first ${FUZZ_FIRST:-5000} --count ${FUZZ_COUNT:-300} 2>&1 | tail -13
  echo "== checked build: smoke()"
  python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
  echo "== ORBFE_DCHECK failures / guard reports in the output above: $(grep -c 'ORBFE_DCHECK failed\|ORBFE_CHECKS:' $OUT/checked_$TAG.log 2>/dev/null || echo 0)"
} > $OUT/checked_$TAG.log 2>&1
cp /tmp/liborbfe_normal.so my_orb_slam2_b200/liborbfe.so
tail -12 $OUT/checked_$TAG.log
