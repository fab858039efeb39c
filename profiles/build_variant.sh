#!/bin/bash
# build_variant.sh <name> "<extra nvcc flags>" [git-rev]: a second build of the library for same-box A/B runs
# (DG1D_LIB=fem_b200/libdg1d_<name>.so python bench.py ...).  Sources are copied to build/<name>/ (git-ignored);
# with a git revision the sources of that revision are used instead of the working tree.
set -e
name=$1; flags=$2; rev=$3
root=$(cd "$(dirname "$0")/.." && pwd)
dst=$root/build/$name
rm -rf "$dst"; mkdir -p "$dst"
if [ -n "$rev" ]; then
  (cd "$root" && git archive "$rev" fem_b200/csrc) | tar -x -C "$dst" --strip-components=2
else
  cp "$root"/fem_b200/csrc/*.cu "$root"/fem_b200/csrc/*.cuh "$root"/fem_b200/csrc/Makefile "$dst"/
fi
make -s -C "$dst" -j8 TUNE="$flags" OUT="$root/fem_b200/libdg1d_$name.so"
ls -la "$root/fem_b200/libdg1d_$name.so"
