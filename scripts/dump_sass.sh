#!/usr/bin/env bash
# Full SASS listing of every kernel of the native library, one file per translation unit, under profiles/sass/
# (instruction encodings stripped).  Run after a build: bash scripts/dump_sass.sh
set -eu
cd "$(dirname "$0")/.."
mkdir -p profiles/sass
for o in attackfl_b200/lib/*.cu.o; do
  name=$(basename "$o" .cu.o)
  cuobjdump -sass "$o" | grep -v -E '^\s*/\* 0x[0-9a-f]+ \*/\s*$' | sed -E 's#\s*/\* 0x[0-9a-f]+ \*/\s*$##' > "profiles/sass/$name.sass"
  echo "$name: $(grep -c 'Function :' profiles/sass/$name.sass) kernels, $(wc -l < profiles/sass/$name.sass) lines"
done
