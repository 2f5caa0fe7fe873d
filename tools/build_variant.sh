#!/bin/bash
# Build a library variant for tools/ab.sh:   bash tools/build_variant.sh NAME [GIT_REF]
#   no GIT_REF: the working tree as it is;  GIT_REF: that commit's csrc/ + include/ (checked out into a temp dir)
set -e
NAME=$1; REF=${2:-}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p $ROOT/tools/scratch/var
if [ -z "$REF" ]; then
  make -C $ROOT/my_orb_slam2_b200/csrc -j8 > /dev/null
  cp $ROOT/my_orb_slam2_b200/liborbfe.so $ROOT/tools/scratch/var/$NAME.so
else
  T=$(mktemp -d)
  git -C $ROOT archive $REF my_orb_slam2_b200/csrc include | tar -x -C $T
  make -C $T/my_orb_slam2_b200/csrc -j8 > /dev/null
  cp $T/my_orb_slam2_b200/liborbfe.so $ROOT/tools/scratch/var/$NAME.so
  rm -rf $T
fi
ls -la $ROOT/tools/scratch/var/$NAME.so
