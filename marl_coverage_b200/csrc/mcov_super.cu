// mcov_super.cu — SuperGridRL kernels (placeholder until the Dec path is parity-green on a B200).
#include "mcov_internal.h"

namespace mcov {
int super_grid_size(const mcov_env* env) { return env->num_sms; }
cudaError_t launch_super(mcov_env*, SuperParams&, cudaStream_t) { return cudaErrorNotSupported; }
}  // namespace mcov
