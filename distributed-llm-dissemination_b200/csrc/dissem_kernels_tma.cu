// placeholder until the TMA bulk-staged variant lands
#include "dissem_kernels.cuh"
namespace dissem {
cudaError_t launch_copy_csum_tma(const XferDesc *, int, long long, int, SchedCounters *, const LaunchCfg &, cudaStream_t, int *) {
    return cudaErrorNotSupported;
}
}  // namespace dissem
