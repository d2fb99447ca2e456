// TEST INFRASTRUCTURE: extern "C" doorway to the reference's own rowwise quantizer in
// third_party/fbgemm/src/QuantUtils.cc (FloatOrHalfToFusedNBitRowwiseQuantizedSBHalfRef :499, the scalar
// definition, and FloatOrHalfToFusedNBitRowwiseQuantizedSBHalf :574, the dispatcher that
// fbgemm_gpu/src/quantize_ops_cpu.cpp calls), which the Makefile compiles UNMODIFIED where they lie, together
// with QuantUtilsAvx2.cc and Utils.cc.  Nothing here restates reference code.
//
// The reference asks its third-party `cpuinfo` library which ISA the host has; oracle/cpuinfo_stub.cc answers
// from the compiler's CPUID builtins, so the dispatcher takes the same AVX2 path on an AVX2 host as a real build.
#include <cstddef>
#include <cstdint>
#include <cstring>

#include <cpuinfo.h>

#include "fbgemm/QuantUtils.h"

extern "C" {

// out rows: [ncols x uint8][fp16 scale][fp16 bias], ncols + 4 bytes each (bit_rate 8).
void ref_quantize_sbhalf_ref(int bit_rate, const float* in, size_t rows, int cols, uint8_t* out) {
  fbgemm::FloatOrHalfToFusedNBitRowwiseQuantizedSBHalfRef<float>(bit_rate, in, rows, cols, out);
}

void ref_quantize_sbhalf(int bit_rate, const float* in, size_t rows, int cols, uint8_t* out) {
  fbgemm::FloatOrHalfToFusedNBitRowwiseQuantizedSBHalf<float>(bit_rate, in, rows, cols, out);
}

int ref_quantize_has_avx2(void) {
  cpuinfo_initialize();
  return cpuinfo_isa.avx2 ? 1 : 0;
}

}  // extern "C"
