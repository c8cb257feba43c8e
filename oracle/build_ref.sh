#!/usr/bin/env bash
# TEST INFRASTRUCTURE. Compiles the UNMODIFIED reference rasterizers from where
# they lie under /root/reference into oracle/_ref/ (git-ignored, but it travels
# to the GPU box with the gpurun snapshot). No reference source is copied into
# the repo: the only sources compiled from this repo are the two extern "C"
# shims (oracle/ref_shim_*.cu).
#
# The single deviation from the reference's own setup.py flags is
# `-include cstdint` (gcc 13 no longer leaks uint32_t / std::uintptr_t into
# rasterizer_impl.h:24,40-61) and the arch (`compute_100` — the reference's
# CMakeLists targets 70;75;86 and its setup.py passes none).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/submodules" ]; then
    echo "[build_ref] $REF not present: keeping prebuilt oracle/_ref as is" >&2
    exit 0
fi
mkdir -p "$OUT/obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
ARCH="-gencode arch=compute_100,code=sm_100"
COMMON="-O3 -std=c++17 -Xcompiler -fPIC -include cstdint $ARCH"

build_one() {   # name srcdir shim files...
    local name="$1" dir="$2" shim="$3"; shift 3
    local objs=()
    for f in "$@"; do
        local o="$OUT/obj/${name}_$(basename "${f%.cu}").o"
        if [ ! -f "$o" ] || [ "$dir/cuda_rasterizer/$f" -nt "$o" ]; then
            $NVCC $COMMON -I"$dir/third_party/glm" -I"$dir/cuda_rasterizer" \
                -c "$dir/cuda_rasterizer/$f" -o "$o" &
        fi
        objs+=("$o")
    done
    local so="$OUT/obj/${name}_shim.o"
    $NVCC $COMMON -I"$dir/third_party/glm" -I"$dir/cuda_rasterizer" -c "$shim" -o "$so" &
    wait
    $NVCC -shared $ARCH -Xlinker -Bsymbolic -o "$OUT/lib${name}.so" "${objs[@]}" "$so"
    echo "[build_ref] built $OUT/lib${name}.so"
}

build_one ref_dgr "$REF/submodules/diff-gaussian-rasterization" "$HERE/ref_shim_dgr.cu" \
    rasterizer_impl.cu forward.cu backward.cu
build_one ref_r2 "$REF/submodules/rasterizer2" "$HERE/ref_shim_r2.cu" \
    rasterizer_impl.cu forward.cu backward.cu raster.cu
