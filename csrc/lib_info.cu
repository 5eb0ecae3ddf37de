// Library bookkeeping: ABI version and a process-wide launch counter (bench.py reports it).
#include "../include/gfedntm_kernels.h"
#include <atomic>

namespace gf {
std::atomic<long long> g_launches{0};
}

extern "C" int gf_abi_version(void) { return 1; }
extern "C" long long gf_launch_count(void) { return gf::g_launches.load(); }

// sizeof() of every argument struct, so the ctypes mirror in gfedntm_b200/ops/_lib.py can be
// checked against the compiled layout at load time.
extern "C" int gf_sizeof(int which) {
  switch (which) {
    case 0: return (int)sizeof(GfSweepArgs);
    case 1: return (int)sizeof(GfEncArgs);
    case 2: return (int)sizeof(GfLatentArgs);
    case 3: return (int)sizeof(GfLatentBwdArgs);
    case 4: return (int)sizeof(GfEncBwdArgs);
    case 5: return (int)sizeof(GfDecSparseArgs);
    case 6: return (int)sizeof(GfFabricRank);
    case 7: return (int)sizeof(GfFedArgs);
    case 8: return (int)sizeof(GfFedRange);
    default: return -1;
  }
}
