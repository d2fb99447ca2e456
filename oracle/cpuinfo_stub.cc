// TEST INFRASTRUCTURE: the reference's FBGEMM sources ask its third-party `cpuinfo` library (a separate C
// library, not built here) which ISA the host has.  The four symbols they use are answered from the compiler's
// CPUID builtins, so the reference's dispatchers / JIT generators pick the same AVX2 or AVX-512 path as a real
// build would on this host.  Nothing here restates reference code.
#include <cstdint>
#include <cstring>

#include <cpuid.h>
#include <cpuinfo.h>

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
    cpuinfo_isa.avx512f = __builtin_cpu_supports("avx512f");
    cpuinfo_isa.avx512bw = __builtin_cpu_supports("avx512bw");
    cpuinfo_isa.avx512dq = __builtin_cpu_supports("avx512dq");
    cpuinfo_isa.avx512vl = __builtin_cpu_supports("avx512vl");
    cpuinfo_isa.avx512vnni = __builtin_cpu_supports("avx512vnni");
    done = true;
  }
  return true;
}

// fbgemmIsIntelXeonD() (Utils.cc, consulted by fbgemmInstructionSet()) looks for "Xeon(R) D-" in the first
// package's name: one package carrying the CPUID brand string.
const struct cpuinfo_package* cpuinfo_get_packages(void) {
  static struct cpuinfo_package pkg;
  static bool done = false;
  if (!done) {
    std::memset(&pkg, 0, sizeof(pkg));
    unsigned int regs[12] = {0};
    unsigned int max_ext = __get_cpuid_max(0x80000000u, nullptr);
    if (max_ext >= 0x80000004u) {
      for (unsigned int i = 0; i < 3; ++i)
        __get_cpuid(0x80000002u + i, &regs[4 * i], &regs[4 * i + 1], &regs[4 * i + 2], &regs[4 * i + 3]);
      std::memcpy(pkg.name, regs, sizeof(regs) < sizeof(pkg.name) - 1 ? sizeof(regs) : sizeof(pkg.name) - 1);
    }
    done = true;
  }
  return &pkg;
}
uint32_t cpuinfo_get_packages_count(void) { return 1; }

}  // extern "C"
