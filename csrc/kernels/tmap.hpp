// Host-side helpers shared by the tcgen05 translation units (defined in gemm_sm100.cu).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace ddpx {

// Cached cuTensorMapEncodeTiled: rank-N tiled map with the 128-byte swizzle (atom32 = the 32-byte-atom flavour MN-major
// tf32 operands need); dims / box innermost first, strides in bytes for dims 1..rank-1.  Returns a DDPX_* code.
int make_tmap(CUtensorMap* out, const void* base, bool bf16, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, bool atom32 = false);

int num_sms();

}  // namespace ddpx
