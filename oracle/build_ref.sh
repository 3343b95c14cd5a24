#!/usr/bin/env bash
# TEST INFRASTRUCTURE.  Builds (1) oracle/_ref/liboracle.so, the plain-C restatement, and
# (2) when /root/reference is present, oracle/_ref/libphmm_ref.so: the reference's own
# pair-HMM sources compiled IN PLACE (nothing is copied) plus oracle/ref_harness.cpp.
# The reference's own build (gradle + cmake) cannot run here: its CMakeLists.txt files were
# never committed (root .gitignore ignores *.txt) and there is no JDK; see SURVEY.md 8c.
#   canonical oracle     : -ffp-contract=off  (AVX-256 and AVX-512 TUs bit-identical)
#   libphmm_ref_fma.so   : g++ defaults (the AVX-512 TU gets contracted vfmadd) -- tolerance check only
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref"
REF="${PHMM_REFERENCE_ROOT:-/root/reference}"
P="$REF/tools/GKL/src/main/native"
mkdir -p "$OUT"

gcc -O2 -std=c11 -fPIC -shared -fopenmp -ffp-contract=off \
    -o "$OUT/liboracle.so" "$HERE/phmm_oracle.c" -lm
echo "built $OUT/liboracle.so"
gcc -O2 -std=c11 -fPIC -shared -fopenmp -o "$OUT/libsw_oracle.so" "$HERE/sw_oracle.c"
echo "built $OUT/libsw_oracle.so"

if [ ! -d "$P/pairhmm" ]; then
  echo "reference sources not found at $P -- keeping any prebuilt $OUT/libphmm_ref.so"
  exit 0
fi

# IntelPairHmmCSource.cpp:7 includes IntelPairHmm.h -> <jni.h>; there is no JDK here.
mkdir -p "$OUT/stub"
cat > "$OUT/stub/jni.h" <<'EOS'
#ifndef PHMM_STUB_JNI_H
#define PHMM_STUB_JNI_H
#define JNIEXPORT
#define JNICALL
typedef void JNIEnv; typedef void* jclass; typedef void* jobject; typedef void* jobjectArray;
typedef void* jdoubleArray; typedef unsigned char jboolean; typedef int jint;
#endif
EOS

build_variant () {   # $1 = output name, $2 = extra flags
  local name="$1" extra="$2" obj="$OUT/obj_$1"
  mkdir -p "$obj"
  local CXXF="-O2 -std=c++11 -fPIC -fopenmp $extra -I$OUT/stub -I$P/common -I$P/pairhmm"
  g++ $CXXF -mavx -c "$P/pairhmm/avx_impl.cc" -o "$obj/avx_impl.o"
  g++ $CXXF -mavx -mavx2 -mavx512f -mavx512dq -mavx512vl -c "$P/pairhmm/avx512_impl.cc" -o "$obj/avx512_impl.o"
  g++ $CXXF -c "$P/pairhmm/pairhmm_common.cc" -o "$obj/pairhmm_common.o"
  g++ $CXXF -c "$P/pairhmm/IntelPairHmmCSource.cpp" -o "$obj/IntelPairHmmCSource.o"
  g++ $CXXF -c "$HERE/ref_harness.cpp" -o "$obj/ref_harness.o"
  g++ -shared -fopenmp -Wl,-Bsymbolic -o "$OUT/$name" "$obj"/*.o -lm
  rm -rf "$obj"
  echo "built $OUT/$name"
}
build_variant libphmm_ref.so "-ffp-contract=off"
build_variant libphmm_ref_fma.so ""

# ---- Smith-Waterman: avx2_impl.cc / avx512_impl.cc include PairWiseSW.h ----
S="$P/smithwaterman"
obj="$OUT/obj_sw"; mkdir -p "$obj"
SWF="-O2 -std=c++11 -fPIC -fopenmp -w -I$P/common -I$S"
g++ $SWF -mavx -mavx2 -c "$S/avx2_impl.cc" -o "$obj/avx2_impl.o"
g++ $SWF -mavx -mavx2 -mavx512f -mavx512dq -mavx512vl -mavx512bw -c "$S/avx512_impl.cc" -o "$obj/avx512_impl.o"
g++ $SWF -c "$HERE/sw_ref_harness.cpp" -o "$obj/sw_ref_harness.o"
g++ -shared -fopenmp -Wl,-Bsymbolic -o "$OUT/libsw_ref.so" "$obj"/*.o
rm -rf "$obj"
echo "built $OUT/libsw_ref.so"

# ---- the reference's own runnable driver, UNMODIFIED, linked once against the reference objects and once
# against libphmm_cuda.so (the link-time drop-in of INTEGRATION.md section 1).  benchmarks/phmm's drivers need
# PALISADE headers; tools/GKL/src/examples/PairHMM/PairHMMUnitTest.cpp is the same parser and the same
# computelikelihoodsboth call without them (SURVEY.md 8c "Drivers").
EX="$REF/tools/GKL/src/examples/PairHMM"
PKG="$(cd "$HERE/.." && pwd)/genomicsbench_palisade_b200"
if [ -f "$EX/PairHMMUnitTest.cpp" ]; then
  g++ -O2 -std=c++11 -fopenmp -w -I"$EX" "$EX/PairHMMUnitTest.cpp" -o "$OUT/gkl_example_driver_ref" \
      -L"$OUT" -lphmm_ref -Wl,-rpath,'$ORIGIN'
  echo "built $OUT/gkl_example_driver_ref"
  if [ -f "$PKG/libphmm_cuda.so" ]; then
    g++ -O2 -std=c++11 -fopenmp -w -I"$EX" "$EX/PairHMMUnitTest.cpp" -o "$OUT/gkl_example_driver_cuda" \
        -L"$PKG" -lphmm_cuda -Wl,-rpath,'$ORIGIN/../../genomicsbench_palisade_b200'
    echo "built $OUT/gkl_example_driver_cuda"
  fi
fi

# ---- the BENCHMARK's own driver, UNMODIFIED (benchmarks/phmm/PairHMMUnitTest_orig.cpp: OpenMP loop over batches,
# one computelikelihoodsboth call per batch and thread :271-299, "Kernel runtime" line :327), for timing the
# link-time drop-in.  Its shacc_pairhmm.h pulls in the fork's palisade_header.h -> <palisade.h>, which is not
# installed; the plaintext driver uses none of it, so a stub header with the few type names palisade_header.h
# mentions (written here, 20 lines, no PALISADE code) is all it needs.  Linked against the reference objects
# (libphmm_ref.so carries IntelPairHmmCSource.cpp as is, fork printf included) and against libphmm_cuda.so.
BP="$REF/benchmarks/phmm"
if [ -f "$BP/PairHMMUnitTest_orig.cpp" ]; then
  PS="$OUT/stub_palisade"
  mkdir -p "$PS/scheme/bfvrns"
  cat > "$PS/palisade.h" <<'STUB'
// Stub (oracle/build_ref.sh): type names only, so that the unmodified plaintext driver compiles without PALISADE.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>
namespace lbcrypto {
class DCRTPoly {};
template <class T> class CiphertextImpl {};
template <class T> using Ciphertext = std::shared_ptr<CiphertextImpl<T>>;
class PlaintextImpl {};
using Plaintext = std::shared_ptr<PlaintextImpl>;
template <class T> class CryptoContextImpl {};
template <class T> using CryptoContext = std::shared_ptr<CryptoContextImpl<T>>;
template <class T> struct LPKeyPair {};
enum SecurityLevel { HEStd_128_classic };
}
STUB
  for f in ciphertext-ser.h cryptocontext-ser.h pubkeylp-ser.h scheme/bfvrns/bfvrns-ser.h; do echo "#pragma once" > "$PS/$f"; done
  g++ -O2 -std=c++11 -fopenmp -w -I"$PS" -I"$BP" "$BP/PairHMMUnitTest_orig.cpp" -o "$OUT/phmm_orig_driver_ref" \
      -L"$OUT" -lphmm_ref -Wl,-rpath,'$ORIGIN'
  echo "built $OUT/phmm_orig_driver_ref"
  if [ -f "$PKG/libphmm_cuda.so" ]; then
    g++ -O2 -std=c++11 -fopenmp -w -I"$PS" -I"$BP" "$BP/PairHMMUnitTest_orig.cpp" -o "$OUT/phmm_orig_driver_cuda" \
        -L"$PKG" -lphmm_cuda -Wl,-rpath,'$ORIGIN/../../genomicsbench_palisade_b200'
    echo "built $OUT/phmm_orig_driver_cuda"
  fi
fi
