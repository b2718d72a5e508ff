#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds oracle/_ref/: the UNMODIFIED nadult/Snail reference hot path,
# compiled from the sources where they lie under $SNAIL_REF (default /root/reference), plus the
# C API in oracle/ref_capi.cpp. Outputs go only into oracle/_ref/ (git-ignored, travels with gpurun).
#
#   oracle/_ref/libsnailref_sse.so     veclib SSE back-end (rcpps/rsqrtps + 1 NR)  -> timed CPU baseline
#   oracle/_ref/libsnailref_scalar.so  veclib scalar back-end (exact 1/x, 1/sqrt)   -> canonical oracle
#   oracle/_ref/libsnailref_depth2.so  the scalar build with `cache.reflections < 1` relaxed to `< 2` (generated copy of
#                                      src/scene_trace.cpp under gen-depth2/): comparison for snl_set_reflection_depth(2)
#   oracle/_ref/libsnailref_b200adapter.so  the scalar build + oracle/ref_adapter.cpp: the product's include/scene_b200.h
#                                      compiled inside the reference, linked with libsnail_b200.so (adapter test only)
#   oracle/_ref/scenes/, oracle/_ref/data/  copies of the reference's scene/texture ASSETS (not sources)
#
# Two generated headers (build artefacts, never committed) make the in-place sources build with g++:
#   gen/rtbase_math.h  = src/rtbase_math.h with `template <> static constexpr` -> `inline constexpr`
#                        (GCC rejects a storage class on an explicit specialisation; upstream uses clang),
#                        force-included so its include guard shadows the original;
#   gen-scalar/veclib.h = veclib/veclib.h with the SSE selector disabled (veclib/veclib.h:16-36,61-80).
# The reference's own build system is not run.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SNAIL_REF:-/root/reference}"
OUT="$HERE/_ref"
JOBS="${JOBS:-$(nproc)}"

if [ ! -d "$REF/src" ]; then
	echo "build_ref: $REF not present; keeping prebuilt oracle/_ref as is" >&2
	exit 0
fi

mkdir -p "$OUT/gen" "$OUT/gen-scalar" "$OUT/scenes" "$OUT/data"

sed 's/template <> static constexpr/template <> inline constexpr/' "$REF/src/rtbase_math.h" > "$OUT/gen/rtbase_math.h.tmp"
cmp -s "$OUT/gen/rtbase_math.h.tmp" "$OUT/gen/rtbase_math.h" 2>/dev/null || mv "$OUT/gen/rtbase_math.h.tmp" "$OUT/gen/rtbase_math.h"
rm -f "$OUT/gen/rtbase_math.h.tmp"

# scalar selector: drop the intrinsics include block and force the #else (scalar) branch
python3 - "$REF/veclib/veclib.h" "$OUT/gen-scalar/veclib.h" <<'EOF'
import re, sys, os
src = open(sys.argv[1]).read()
src = src.replace('#include "veclib_conf.h"', '#include "veclib_conf.h"\n#undef VECLIB_SSE\n#define VECLIB_FORCE_SCALAR 1', 1)
# first selector: intrinsics headers
src = src.replace('#else\n\t\t#error "undefined"\n#endif', '#else\n#endif', 1)
new = src
old = open(sys.argv[2]).read() if os.path.exists(sys.argv[2]) else None
if old != new:
    open(sys.argv[2], 'w').write(new)
EOF

# depth-2 variant (BASELINE config 5 "reflection depth 2" is beyond the reference's own cap): the ONE condition
# `cache.reflections < 1` (src/scene_trace.cpp:454) relaxed to `< 2` in a generated copy of that translation unit
mkdir -p "$OUT/gen-depth2"
sed 's/cache\.reflections < 1/cache.reflections < 2/' "$REF/src/scene_trace.cpp" > "$OUT/gen-depth2/scene_trace.cpp.tmp"
grep -q 'cache.reflections < 2' "$OUT/gen-depth2/scene_trace.cpp.tmp" || { echo "build_ref: depth-2 patch did not apply" >&2; exit 1; }
cmp -s "$OUT/gen-depth2/scene_trace.cpp.tmp" "$OUT/gen-depth2/scene_trace.cpp" 2>/dev/null || mv "$OUT/gen-depth2/scene_trace.cpp.tmp" "$OUT/gen-depth2/scene_trace.cpp"
rm -f "$OUT/gen-depth2/scene_trace.cpp.tmp"

cp -u "$REF"/scenes/*.obj "$REF"/scenes/*.mtl "$OUT/scenes/" 2>/dev/null || true
cp -u "$REF"/data/*.png "$OUT/data/" 2>/dev/null || true

SRC_TUS="triangle bounding_box ray_group rtbase tree_stats base_scene thread_pool camera
	bvh/tree bvh/traverse dbvh/tree dbvh/traverse ray_generator scene scene_trace render
	formats/wavefront_obj shading/material shading/uber_material
	sampling/point_sampler sampling/sampler mipmap_texture"
FWK_TUS="base_vector format str sys_base enum parse type_info logger any
	sys/assert sys/assert_impl sys/error sys/exception sys/expected sys/memory sys/on_fail sys/thread
	io/file_stream io/file_system io/stream io/memory_stream
	gfx/image gfx/image_stbi gfx/image_tga gfx/color io/xml"

BASE="-std=c++2a -O3 -msse2 -ffp-contract=off -fPIC -fpermissive -w -fno-strict-aliasing
	-I$REF/libfwk/include -I$REF/libfwk/src -I$REF/libfwk/extern
	-DFWK_PLATFORM_LINUX -DFWK_DWARF_DISABLED -DNDEBUG"
COMMON="$BASE -I$REF/src -include $OUT/gen/rtbase_math.h -D'ASSERT_EQ(a,b)=ASSERT((a)==(b))'"
FWKFLAGS="$BASE -DFATAL=FWK_FATAL"

# the adapter library needs the product library to link against
ADAPTER=""
[ -f "$HERE/../snail_b200/libsnail_b200.so" ] && ADAPTER="$OUT/libsnailref_b200adapter.so"

MK="$OUT/Makefile.gen"
{
	echo "CXX=g++"
	echo "COMMON=$(echo $COMMON)"
	echo "FWKFLAGS=$(echo $FWKFLAGS)"
	echo "all: $OUT/libsnailref_sse.so $OUT/libsnailref_scalar.so $OUT/libsnailref_depth2.so $ADAPTER"
	FWK_OBJS=""
	for tu in $FWK_TUS; do
		o="$OUT/obj/fwk/$(echo $tu | tr / _).o"
		FWK_OBJS="$FWK_OBJS $o"
		echo "$o: $REF/libfwk/src/$tu.cpp"
		echo "	@mkdir -p \$(dir \$@)"
		echo "	\$(CXX) \$(FWKFLAGS) -c \$< -o \$@"
	done
	for variant in sse scalar; do
		if [ $variant = sse ]; then VINC="-I$REF/veclib"; else VINC="-I$OUT/gen-scalar -I$REF/veclib"; fi
		OBJS=""
		for tu in $SRC_TUS; do
			o="$OUT/obj/$variant/$(echo $tu | tr / _).o"
			OBJS="$OBJS $o"
			echo "$o: $REF/src/$tu.cpp $OUT/gen/rtbase_math.h $OUT/gen-scalar/veclib.h"
			echo "	@mkdir -p \$(dir \$@)"
			echo "	\$(CXX) \$(COMMON) $VINC -c \$< -o \$@"
		done
		for tu in ref_capi ref_shim; do
			o="$OUT/obj/$variant/$tu.o"
			OBJS="$OBJS $o"
			echo "$o: $HERE/$tu.cpp $OUT/gen/rtbase_math.h $OUT/gen-scalar/veclib.h"
			echo "	@mkdir -p \$(dir \$@)"
			echo "	\$(CXX) \$(COMMON) $VINC -c \$< -o \$@"
		done
		echo "$OUT/libsnailref_$variant.so: $OBJS $FWK_OBJS"
		echo "	\$(CXX) -shared -o \$@ \$^ -lpthread -Wl,--no-undefined"
		if [ $variant = scalar ]; then
			# the scalar build with scene_trace.cpp replaced by its depth-2 copy
			o="$OUT/obj/depth2/scene_trace.o"
			echo "$o: $OUT/gen-depth2/scene_trace.cpp $OUT/gen/rtbase_math.h $OUT/gen-scalar/veclib.h"
			echo "	@mkdir -p \$(dir \$@)"
			echo "	\$(CXX) \$(COMMON) $VINC -c \$< -o \$@"
			echo "$OUT/libsnailref_depth2.so: $o $(echo $OBJS | tr ' ' '\n' | grep -v '/scene_trace.o$' | tr '\n' ' ') $FWK_OBJS"
			echo "	\$(CXX) -shared -o \$@ \$^ -lpthread -Wl,--no-undefined"
		fi
		if [ $variant = scalar ] && [ -n "$ADAPTER" ]; then
			# the product's reference-side adapter (include/scene_b200.h) compiled inside the reference and linked with
			# libsnail_b200.so into a library of its own; render.o is left out: ref_adapter.cpp includes src/render.cpp
			o="$OUT/obj/$variant/ref_adapter.o"
			echo "$o: $HERE/ref_adapter.cpp $HERE/../include/scene_b200.h $HERE/../include/snail_b200.h $REF/src/render.cpp $OUT/gen/rtbase_math.h"
			echo "	@mkdir -p \$(dir \$@)"
			echo "	\$(CXX) \$(COMMON) $VINC -I$HERE/../include -c \$< -o \$@"
			echo "$ADAPTER: $o $(echo $OBJS | tr ' ' '\n' | grep -v '/render.o$' | tr '\n' ' ') $FWK_OBJS $HERE/../snail_b200/libsnail_b200.so"
			echo "	\$(CXX) -shared -o \$@ $o $(echo $OBJS | tr ' ' '\n' | grep -v '/render.o$' | tr '\n' ' ') $FWK_OBJS -L$HERE/../snail_b200 -lsnail_b200 -Wl,-rpath,'\$\$ORIGIN/../../snail_b200' -lpthread -Wl,--no-undefined"
		fi
	done
} > "$MK"

make -s -j"$JOBS" -f "$MK" all
echo "build_ref: ok -> $OUT/libsnailref_sse.so $OUT/libsnailref_scalar.so"
