// trace_inst.cu -- compiled once per step method with -DRTB_METHOD=K (K = 1..11) so the
// eleven kernels build in parallel and each object holds only its own advancer/updater.
#include "trace_kernel.cuh"

#ifndef RTB_METHOD
#error "compile with -DRTB_METHOD=<1..11>"
#endif

#define RTB_CAT2(a, b) a##b
#define RTB_CAT(a, b) RTB_CAT2(a, b)

namespace rtb {
cudaError_t RTB_CAT(launch_trace_m, RTB_METHOD)(const TraceArgs &a, const LaunchCfg &c, cudaStream_t st)
{
    return launch_trace<RTB_METHOD>(a, c, st);
}
} // namespace rtb
