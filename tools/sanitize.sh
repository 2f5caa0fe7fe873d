#!/bin/bash
# compute-sanitizer pass on the GPU box:   bash tools/sanitize.sh r2
# memcheck + racecheck + initcheck + synccheck over smoke() (EuRoC pair: extract L/R, stereo, brute force) and over the
# odd-shape extractor tests (TMA edge tiles, ragged cells).  Logs go to gpurun_out/sanitizer_<tool>_<what>_<tag>.log;
# tools/make_sanitizer_summary.py turns them into profiles/<tag>_sanitizer.md.
set -u
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
CS=${CS:-compute-sanitizer}
# the sanitizer needs the non-graph launch path to attribute reports to kernels; ORBFE_NO_GRAPHS=1 selects it
export ORBFE_NO_GRAPHS=1
for TOOL in memcheck racecheck initcheck synccheck; do
  EXTRA=""
  [ $TOOL = initcheck ] && EXTRA="--track-unused-memory no"
  timeout 900 $CS --tool $TOOL $EXTRA --print-limit 40 --log-file $OUT/sanitizer_${TOOL}_smoke_$TAG.log \
      python -c "import __graft_entry__ as g; g.smoke()" > $OUT/sanitizer_${TOOL}_smoke_$TAG.out 2>&1
  echo "$TOOL smoke rc=$?"
  timeout 900 $CS --tool $TOOL $EXTRA --print-limit 40 --log-file $OUT/sanitizer_${TOOL}_odd_$TAG.log \
      python -m pytest tests/test_gpu_extractor.py -q -x -m gpu -k "odd_shapes or stress" > $OUT/sanitizer_${TOOL}_odd_$TAG.out 2>&1
  echo "$TOOL odd rc=$?"
  timeout 900 $CS --tool $TOOL $EXTRA --print-limit 40 --log-file $OUT/sanitizer_${TOOL}_match_$TAG.log \
      python -m pytest tests/test_gpu_matcher.py tests/test_gpu_stereo.py -q -x -m gpu -k "not full and not million" > $OUT/sanitizer_${TOOL}_match_$TAG.out 2>&1
  echo "$TOOL match rc=$?"
done
grep -H "ERROR SUMMARY\|RACECHECK SUMMARY" $OUT/sanitizer_*_$TAG.log
