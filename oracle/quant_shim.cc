// TEST INFRASTRUCTURE: extern "C" doorway to the reference's own rowwise quantizer in
// third_party/fbgemm/src/QuantUtils.cc (FloatOrHalfToFusedNBitRowwiseQuantizedSBHalfRef :499, the scalar
// definition, and FloatOrHalfToFusedNBitRowwiseQuantizedSBHalf :574, the dispatcher that
// fbgemm_gpu/src/quantize_ops_cpu.cpp calls), which the Makefile compiles UNMODIFIED where they lie, together
// with QuantUtilsAvx2.cc and Utils.cc.  Nothing here restates reference code.
//
// The reference asks its third-party `cpuinfo` library which ISA the host has.  cpuinfo is a separate C
// library that is not built here; the two symbols the quantizer needs are answered below from the compiler's
// own CPUID builtin, so the dispatcher takes the same AVX2 path on an AVX2 host as a real build would.
#include <cstddef>
#include <cstdint>
#include <cstring>

#include <cpuinfo.h>

#include "fbgemm/QuantUtils.h"

extern "C" {

struct cpuinfo_x86_isa cpuinfo_isa;

bool cpuinfo_initialize(void) {
  static bool done = false;
  if (!done) {
    std::memset(&cpuinfo_isa, 0, sizeof(cpuinfo_isa));
    __builtin_cpu_init();
    cpuinfo_isa.avx = __builtin_cpu_supports("avx");
    cpuinfo_isa.avx2 = __builtin_cpu_supports("avx2");
    cpuinfo_isa.fma3 = __builtin_cpu_supports("fma");
    cpuinfo_isa.f16c = __builtin_cpu_supports("f16c");
    done = true;  // AVX-512 is left off: the reference's quantizer has no AVX-512 variant
  }
  return true;
}

// Referenced only by fbgemmIsIntelXeonD() in Utils.cc, which the quantizer never calls.
const struct cpuinfo_package* cpuinfo_get_packages(void) { return nullptr; }
uint32_t cpuinfo_get_packages_count(void) { return 0; }

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
