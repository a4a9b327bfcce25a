// placeholder until the BC7 encoder lands
#include "hkbcn_kernels.h"
namespace hkbcn {
uint32_t bc7_tile_blocks() { return 32; }
cudaError_t launch_bc7_encode(const EncodeParams&, int, cudaStream_t) { return cudaErrorNotSupported; }
cudaError_t bc7_upload_tables() { return cudaErrorNotSupported; }
}
