// Row-shard index arithmetic shared by the CSMain and scene kernels and their launchers (ptb_set_row_shard_bands).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ptb {

// Row shard (multi-GPU image tiles): the image is cut into bands of `block` rows; a context owns the bands
// g with g % stride == offset.  shard_row maps the i-th row of the shard to its image row (which may be
// >= height in the ragged last band only if i >= shard_row_count), shard_row_count is how many rows it owns.
__host__ __device__ inline uint32_t shard_row(uint32_t offset, uint32_t stride, uint32_t block, uint32_t i)
{
    return block == 1u ? offset + i * stride : ((i / block) * stride + offset) * block + i % block;
}
inline uint32_t shard_row_count(uint32_t height, uint32_t offset, uint32_t stride, uint32_t block)
{
    const uint32_t bands = (height + block - 1u) / block;
    if (offset >= bands) return 0u;
    const uint32_t mine = (bands - offset + stride - 1u) / stride;
    const uint32_t last = offset + (mine - 1u) * stride;
    return mine * block - (last == bands - 1u ? bands * block - height : 0u);
}

}  // namespace ptb
