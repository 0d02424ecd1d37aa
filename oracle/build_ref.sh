#!/usr/bin/env bash
# TEST INFRASTRUCTURE. Builds oracle/_ref/libsunsky_ref.so: the reference's own sunsky.cpp + sunsky.h +
# Dr.Jit/Mitsuba header-only math, compiled IN PLACE from $SSK_REFERENCE (default /root/reference).
# Nothing from the reference is copied into the repository; generated files stay in oracle/_ref/gen/.
# Does NOT use the reference's build system (cmake); only its resources/configure.py to emit config.h.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SSK_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
GEN="$OUT/gen"
if [ ! -f "$REF/src/emitters/sunsky.cpp" ]; then
    echo "build_ref: reference tree not found at $REF; keeping any prebuilt $OUT/libsunsky_ref.so" >&2
    exit 0
fi
mkdir -p "$GEN"
# 1. config.h for two scalar variants (the macro set is all the headers need)
if [ ! -f "$GEN/include/mitsuba/core/config.h" ]; then
    ( cd "$GEN" &&
      sed 's/@INSERT_CONFIGURATIONS_HERE@/"scalar_rgb", "scalar_spectral"/' "$REF/resources/mitsuba.conf.template" > mitsuba.conf &&
      python3 "$REF/resources/configure.py" GNU 13.3 "$REF" > /dev/null )
fi
# 2. CIE 1931 table block, lifted at build time from the reference's src/core/spectrum.cpp
awk '/^static const Float cie1931_tbl\[/ {p=1} p {print} p && /};/ {exit}' "$REF/src/core/spectrum.cpp" > "$GEN/cie1931_tbl.inc"
test -s "$GEN/cie1931_tbl.inc"
# 3. compile. -mfma + -ffp-contract=off: dr::fmadd is a hardware FMA, a*b+c is not contracted
#    (what the LLVM backend emits: llvm.fma + plain fmul/fadd, no fast-math flags).
CXX="${SSK_CXX:-/usr/bin/g++}"   # the system compiler: links libstdc++ dynamically (a static copy clashes with numpy/torch's)
# -fno-strict-aliasing: Dr.Jit requires it (ext/drjit/CMakeLists.txt:71 makes it an INTERFACE option of the drjit target);
#    without it g++ -O2 miscompiles the scalar instantiation (the sky term of eval() comes out as 0)
FLAGS="-std=c++17 -O3 -fPIC -shared -mfma -ffp-contract=off -fno-fast-math -fno-strict-aliasing -w -pthread"
INC="-I $HERE/ref_shadow -I $GEN -I $GEN/include -I $REF -I $REF/include -I $REF/ext/drjit/include \
 -I $REF/ext/drjit/ext/drjit-core/include -I $REF/ext/drjit/ext/drjit-core/ext/nanothread/include \
 -I $REF/ext/tinyformat -I $REF/ext/fastfloat/include -I $REF/ext/nanobind/include"
# Two ISA builds of the same harness (the library travels prebuilt to a GPU box whose host CPU is not known here):
#   libsunsky_ref.so         AVX2 + FMA: runs anywhere the image runs; the PARITY oracle (flavours 0 and 1) and the
#                            packet baseline on 8-wide packets
#   libsunsky_ref_avx512.so  + AVX-512 F/VL/DQ/BW/CD: 16-wide packets for the timed CPU baseline; loaded only when
#                            /proc/cpuinfo lists those features (oracle/ref_oracle.py)
$CXX $FLAGS -mavx2 $INC "$HERE/ref_harness.cpp" -o "$OUT/libsunsky_ref.so" &
P1=$!
$CXX $FLAGS -mavx2 -mavx512f -mavx512vl -mavx512dq -mavx512bw -mavx512cd $INC "$HERE/ref_harness.cpp" -o "$OUT/libsunsky_ref_avx512.so" &
P2=$!
# second oracle: the original fp64 Hosek-Wilkie model (mi.hosek_sun_rad), see oracle/hosek_harness.cpp
$CXX -std=c++17 -O2 -fPIC -shared -mavx2 -mfma -ffp-contract=off -fno-strict-aliasing -w \
    -I $REF -I $REF/ext/drjit/include -I $REF/ext/drjit/ext/drjit-core/include "$HERE/hosek_harness.cpp" -o "$OUT/libhosek_ref.so" &
P3=$!
wait $P1; wait $P2; wait $P3
echo "built $OUT/libsunsky_ref.so $OUT/libsunsky_ref_avx512.so"
