// SampleSeededOp / SampleOp on the GPU: AES-128-CTR
// keystream written straight into the tensor.  Integer-ALU / shared-memory
// bound (T-table AES), not HBM bound.
#include "vecops.cuh"

namespace {

constexpr int TPB = AES_TPB_LIGHT;

#define AES_LAUNCH(ctx, kern, grid, ...)                                                                        \
  do {                                                                                                          \
    MB_CUDA((ctx), cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AES_SMEM_BYTES));    \
    kern<<<(grid).grid_dim, (grid).block_dim, AES_SMEM_BYTES, (ctx)->stream>>>(__VA_ARGS__);                              \
  } while (0)

template <int K>
__global__ void __launch_bounds__(TPB) sample_kernel(AesKey key, vec_t *__restrict__ out, int64_t nvec, int64_t numel, uint64_t first_block) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane sTl = aes_lane(sT);
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
    vec_t r = prf_vec<K>(sTl, key, first_block + (uint64_t)v);
    if (K == VK_BIT) r = bit_tail_mask(r, v, numel);
    out[v] = r;
  }
}

// Ring kernels with max_value = Some(1): element e = get_bit() = keystream bit e
// (host/ops.rs:1971-1974,2001-2004).  One thread expands one AES block into 128
// ring elements; a rarely used op, kept simple.
template <typename T>
__global__ void __launch_bounds__(TPB) sample_bits_as_ring_kernel(AesKey key, T *__restrict__ out, int64_t numel) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane sTl = aes_lane(sT);
  int64_t nblocks = (numel + 127) / 128;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblocks; b += (int64_t)gridDim.x * blockDim.x) {
    uint32_t o[4];
    aes_ctr_block(sTl, key, (uint64_t)b, o);
#pragma unroll 1
    for (int j = 0; j < 128; j++) {
      int64_t e = b * 128 + j;
      if (e < numel) out[e] = (T)((o[j >> 5] >> (j & 31)) & 1u);
    }
  }
}

}  // namespace

int k_sample(mb_ctx *ctx, const AesKey &key, mb_tensor *out, bool bits_as_ring, uint64_t first_block) {
  if (out->numel == 0) return MB_OK;
  if (bits_as_ring) {
    int64_t nblocks = (out->numel + 127) / 128;
    const mb_aes_cfg_t g = mb_aes_cfg(ctx, nblocks, TPB);
    if (out->dtype == MB_RING128)
      AES_LAUNCH(ctx, sample_bits_as_ring_kernel<u128>, g, key, (u128 *)out->dptr, out->numel);
    else
      AES_LAUNCH(ctx, sample_bits_as_ring_kernel<uint64_t>, g, key, (uint64_t *)out->dptr, out->numel);
    MB_LAUNCH_CHECK_W(ctx, "sample_bits_as_ring", (double)out->nbytes, (double)nblocks, 0);
    return MB_OK;
  }
  int64_t nvec = vec_count(out);
  const mb_aes_cfg_t g = mb_aes_cfg(ctx, nvec, TPB);
  switch (vec_kind(out->dtype)) {
    case VK_RING128: AES_LAUNCH(ctx, sample_kernel<VK_RING128>, g, key, (vec_t *)out->dptr, nvec, out->numel, first_block); break;
    case VK_RING64: AES_LAUNCH(ctx, sample_kernel<VK_RING64>, g, key, (vec_t *)out->dptr, nvec, out->numel, first_block); break;
    case VK_BIT: AES_LAUNCH(ctx, sample_kernel<VK_BIT>, g, key, (vec_t *)out->dptr, nvec, out->numel, first_block); break;
  }
  MB_LAUNCH_CHECK_W(ctx, "sample", 16.0 * nvec, (double)nvec, 0);
  return MB_OK;
}
