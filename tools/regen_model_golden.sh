#!/bin/bash
# Regenerates tests/golden/model_golden.json: the reference's own Model.cpp (compiled in place by oracle/build_ref.sh)
# loads the staged Sponza (real glTF + textures, stand-in Sponza.bin) and prints sizes and FNV-1a hashes.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
mkdir -p "$ROOT/oracle/_ref/models"
ln -sfn "$ROOT/assets/_ref/sponza" "$ROOT/oracle/_ref/models/sponza"
"$ROOT/oracle/_ref/ref_model_dump" | sed -n '2,$p' > "$ROOT/tests/golden/model_golden.json"   # line 1 is the loader's own log line
rm "$ROOT/oracle/_ref/models/sponza"; rmdir "$ROOT/oracle/_ref/models"
echo "wrote tests/golden/model_golden.json"
