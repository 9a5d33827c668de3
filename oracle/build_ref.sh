#!/bin/bash
# oracle/build_ref.sh -- TEST INFRASTRUCTURE.
# Builds the patched reference from the sources where they lie
# under $RT_REF (default /root/reference). Outputs ONLY binaries into oracle/_ref/:
#   raytrace_ref      the reference CLI (C/Makefile:4 flags + -ffp-contract=off), patches P1+P2
#   raytrace_asis     the as-shipped CLI (P1 only, so it does not crash) -- reported, not graded
#   libref.so         row-range harness (oracle/ref_harness.c) for tests and the CPU baseline
# No reference source is copied into the repo: the patched raycast.c lives in a mktemp dir.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${RT_REF:-/root/reference}/C"
OUT="$HERE/_ref"
if [ ! -f "$REF/raycast.c" ]; then
  echo "build_ref.sh: $REF/raycast.c not found; keeping prebuilt oracle/_ref" >&2
  exit 0
fi
mkdir -p "$OUT"
TMP="$(mktemp -d /tmp/rtref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
# P1: zero num_lights (C/raycast.c:44).  P2: reset next_nearest_object (before C/raycast.c:357).
sed -e 's|^  json_struct->num_shapes = 0;$|  json_struct->num_shapes = 0; json_struct->num_lights = 0;|' \
    -e 's|^      next_nearest_object_index = find_nearest_object(json_struct, &next_nearest_object, next_ray_orgin, $|      next_nearest_object = NULL;\n&|' \
    "$REF/raycast.c" > "$TMP/raycast_patched.c"
sed -e 's|^  json_struct->num_shapes = 0;$|  json_struct->num_shapes = 0; json_struct->num_lights = 0;|' \
    "$REF/raycast.c" > "$TMP/raycast_p1.c"
grep -q 'next_nearest_object = NULL;' "$TMP/raycast_patched.c" || { echo "P2 patch did not apply" >&2; exit 1; }
grep -q 'num_lights = 0;' "$TMP/raycast_patched.c" || { echo "P1 patch did not apply" >&2; exit 1; }
CFLAGS="-O3 -ffp-contract=off -w -I$REF"
HFLAGS="$CFLAGS -DNDEBUG"   # harness libraries: shoot()'s assert (C/raycast.c:245) off, see ref_harness.c
SUPPORT="$REF/v3math.c $REF/ppm.c $REF/parse.c $REF/objects.c"
gcc $CFLAGS "$TMP/raycast_patched.c" $SUPPORT -o "$OUT/raytrace_ref" -lm
gcc $CFLAGS "$TMP/raycast_p1.c" $SUPPORT -o "$OUT/raytrace_asis" -lm
gcc $HFLAGS -fPIC -shared -DREF_RAYCAST_C="\"$TMP/raycast_patched.c\"" \
    "$HERE/ref_harness.c" $SUPPORT -o "$OUT/libref.so" -lm -lpthread
# Ray-counting variant (bench.py's reference arm counts the rays of its sample with the reference itself, untimed):
# the call sites of find_nearest_object -- the only places where "find_nearest_object(json_struct," occurs -- go
# through a counting macro of ref_harness.c.  Never used for timing.
sed -e 's|find_nearest_object(json_struct,|ref_counted_fno(json_struct,|' "$TMP/raycast_patched.c" > "$TMP/raycast_count.c"
[ "$(grep -c 'ref_counted_fno(json_struct,' "$TMP/raycast_count.c")" -ge 3 ] || { echo "count patch did not apply" >&2; exit 1; }
gcc $CFLAGS -fPIC -shared -DREF_COUNT_RAYS -DREF_RAYCAST_C="\"$TMP/raycast_count.c\"" \
    "$HERE/ref_harness.c" $SUPPORT -o "$OUT/libref_count.so" -lm -lpthread
# Deep-bounce variant of the same harness: MAX_RECURSION 7 -> 50, the bound the reference's own CUDA and OptiX
# variants use (CUDA/raycast.cu:13 MAX_ITER, Optix/src/deviceCode.cu:5); pins rt_set_max_depth(50).
sed -e 's|^const int MAX_RECURSION = 7;$|const int MAX_RECURSION = 50;|' "$TMP/raycast_patched.c" > "$TMP/raycast_d50.c"
grep -q 'MAX_RECURSION = 50;' "$TMP/raycast_d50.c" || { echo "depth patch did not apply" >&2; exit 1; }
gcc $CFLAGS -fPIC -shared -DREF_RAYCAST_C="\"$TMP/raycast_d50.c\"" \
    "$HERE/ref_harness.c" $SUPPORT -o "$OUT/libref_d50.so" -lm -lpthread
echo "built: $(ls "$OUT")"
# Context comparator: the reference's own CUDA variant, unmodified, rebuilt for
# sm_100a (its Makefile targets sm_75).  Timed by tools/gpu_refcuda.sh on C1-C3 only: its shared-memory
# scene loader and MAX_SHAPES 128 (CUDA/raycast.cu:14,123) rule out C4/C5.
CU="${RT_REF:-/root/reference}/CUDA"
if [ -f "$CU/raycast.cu" ] && command -v nvcc >/dev/null 2>&1; then
  (cd "$CU" && nvcc -O3 -rdc=true -w -gencode=arch=compute_100a,code=sm_100a raycast.cu v3math.cu ppm.cu parse.cu objects.cu \
      -o "$OUT/raytrace_refcuda" --objdir-as-tempdir 2>/dev/null) || \
  (cd "$TMP" && nvcc -O3 -rdc=true -w -gencode=arch=compute_100a,code=sm_100a "$CU/raycast.cu" "$CU/v3math.cu" "$CU/ppm.cu" "$CU/parse.cu" "$CU/objects.cu" \
      -I"$CU" -o "$OUT/raytrace_refcuda") || echo "build_ref.sh: reference CUDA variant did not build (context only)" >&2
fi
echo "built: $(ls "$OUT" | tr '\n' ' ')"
