// dgemm_tc.cu -- exact int8-slice tcgen05 kernel for dgemm_compressed (mode 1).
// Placeholder until the kernel lands: reports "not available" so dgemm_device uses mode 0.
#include "common.cuh"
#include "handle.h"

namespace mb200 {

bool dgemm_tc_available() { return false; }

int dgemm_tc_device(Handle *, int, int, const double *, long, double *, long) {
  return fail(-20, __func__, __LINE__, "tensor-core dgemm_compressed not built");
}

}  // namespace mb200
