// mcov_extras.cu — gated DecGridRL observation layers (distance map, dijkstra path, mini-map).
#include "mcov_internal.h"

namespace mcov {
cudaError_t launch_dec_extras(mcov_env*, float*, int, cudaStream_t) { return cudaErrorNotSupported; }
}  // namespace mcov
