#!/bin/bash
# Compiles the few reference sources that build without Vulkan (Camera.cpp, Model.cpp, Utils.cpp + vendored glm,
# tinygltf) IN PLACE from /root/reference into oracle/_ref/ (git-ignored).  Nothing is copied into the repo.
# The reference's own build system (CMake + Vulkan SDK + glslc) is not run: it cannot work on this box.
set -e
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
    echo "build_ref.sh: $REF not present; keeping prebuilt oracle/_ref (if any)"
    exit 0
fi
mkdir -p "$OUT"
CXX=${CXX:-g++}
FLAGS="-O2 -std=c++17 -ffp-contract=off -w -I$REF/src -I$REF/submodules/glm -DMODELS_FOLDER=\"$OUT/models/\""
$CXX $FLAGS -o "$OUT/ref_camera_dump" "$HERE/ref_driver/ref_camera_dump.cpp" "$REF/src/Camera.cpp"
echo "built $OUT/ref_camera_dump"
if [ -f "$HERE/ref_driver/ref_model_dump.cpp" ]; then
    if [ ! -f "$OUT/tiny_gltf.o" ]; then
        $CXX -O1 -std=c++17 -w -c -I$REF/submodules/tinygltf -o "$OUT/tiny_gltf.o" "$REF/submodules/tinygltf/tiny_gltf.cc"
    fi
    $CXX $FLAGS -I$REF/submodules/tinygltf -DTINYGLTF_NO_INCLUDE_STB_IMAGE_WRITE -o "$OUT/ref_model_dump" \
        "$HERE/ref_driver/ref_model_dump.cpp" "$REF/src/Model.cpp" "$REF/src/Utils.cpp" "$OUT/tiny_gltf.o" 2>&1 | tail -5 || true
    [ -f "$OUT/ref_model_dump" ] && echo "built $OUT/ref_model_dump"
fi
