// bdpt_kernels.cu -- bidirectional path tracer stages (placeholder until the BDPT kernels land:
// the launcher reports cudaErrorNotSupported, so ptb_bdpt_render_sample fails loudly).
#include "bdpt_kernels.h"

namespace mcrt {

const std::vector<const char*>& bdpt_buffer_names() {
    static const std::vector<const char*> names;
    return names;
}
std::vector<BdptBufferSpec> bdpt_buffer_specs(BdptState&, int) { return {}; }

int launch_bdpt_sample(const PtState&, const BdptState&, int, int, cudaStream_t, cudaError_t* err) {
    *err = cudaErrorNotSupported;
    return 0;
}

}  // namespace mcrt
