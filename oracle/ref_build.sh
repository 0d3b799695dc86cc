#!/bin/sh
# Compiles, from the sources where they lie under the reference checkout ($1), the pieces of
# the real reference that build from their own few files:
#   _ref/libref_camera.so  src/camera.cpp (unmodified) + oracle/glm_shim + ref_camera_api.cpp
#   _ref/ref_pack          include/tiny_obj_loader.h driving the packing rules of main.cpp
# The render path itself (main.cpp + GLSL shaders) needs Vulkan, GLFW and GLM: unbuildable
# here.  Outputs go only into oracle/_ref/ (git-ignored, travels with gpurun).
set -e
REF=${1:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
mkdir -p "$HERE/_ref"
g++ -O2 -std=c++17 -fPIC -shared -w -I "$HERE/glm_shim" -I "$REF/include" \
    -o "$HERE/_ref/libref_camera.so" "$REF/src/camera.cpp" "$HERE/ref_camera_api.cpp"
g++ -O2 -std=c++17 -w -I "$REF/include" -o "$HERE/_ref/ref_pack" "$HERE/ref_pack.cpp"
echo "oracle/_ref built from $REF"
