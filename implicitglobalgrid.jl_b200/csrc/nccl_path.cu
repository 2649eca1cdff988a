// NCCL comparison path (placeholder until the grouped ncclSend/ncclRecv variant lands).
#include <cuda_runtime.h>

#include "igg_internal.h"

struct igg_ctx;

namespace igg {
int nccl_update_halo(igg_ctx*, int, const igg_field*, int, const int*, int, cudaStream_t) {
    return fail(IGG_ERR_UNSUPPORTED, "NCCL comparison path not built yet.");
}
void nccl_destroy(igg_ctx*) {}
}  // namespace igg

extern "C" {
size_t igg_nccl_unique_id_size(void) { return 128; }
int igg_nccl_get_unique_id(void*) { return igg::fail(IGG_ERR_UNSUPPORTED, "NCCL comparison path not built yet."); }
int igg_nccl_init(igg_ctx*, const void*) { return igg::fail(IGG_ERR_UNSUPPORTED, "NCCL comparison path not built yet."); }
}
