// PRF microbenchmark: AES-128-CTR blocks per second of the device formulation in moose_b200/csrc/aes.cuh.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 [-DAES_ABS_TABLE=1 -DTPB=1024] -o aes_bench tools/aes_bench.cu moose_b200/csrc/hostcrypto.cpp ...
#include <cstdio>
#include <cstdlib>
#include "../moose_b200/csrc/aes.cuh"
#ifndef TPB
#define TPB 256
#endif
#ifndef NB
#define NB 1
#endif
// NB independent AES chains in flight per thread (vectors threadIdx + j * TPB of the CTA's chunk): the T-table rounds are
// dependent shared-memory lookups, so what one thread cannot overlap the other warps of the SM must.
__global__ void __launch_bounds__(TPB) prf_kernel(AesKey key, uint4 *__restrict__ out, int64_t nvec) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane a = aes_lane(sT);
  for (int64_t base = (int64_t)blockIdx.x * TPB * NB; base < nvec; base += (int64_t)gridDim.x * TPB * NB) {
    uint32_t o[NB][4];
#pragma unroll
    for (int j = 0; j < NB; j++) aes_ctr_block(a, key, (uint64_t)(base + threadIdx.x + j * TPB), o[j]);
#pragma unroll
    for (int j = 0; j < NB; j++) {
      const int64_t v = base + threadIdx.x + j * TPB;
      if (v < nvec) out[v] = make_uint4(o[j][0], o[j][1], o[j][2], o[j][3]);
    }
  }
}
int main(int argc, char **argv) {
  const int64_t nvec = 1ll << 28;  // 4 GB of keystream
  uint4 *out;
  cudaMalloc(&out, nvec * 16);
  AesKey key;
  uint8_t k[16] = {0};
  mb_aes_expand(k, &key);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
#if AES_ABS_TABLE
  const int smem = 0x20000 - 1024, grid = sms;
#else
  const int smem = AES_SMEM_BYTES, grid = sms * (768 / TPB);
#endif
  cudaFuncSetAttribute(prf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int it = 0; it < 3; it++) prf_kernel<<<grid, TPB, smem>>>(key, out, nvec);
  cudaEventRecord(e0);
  for (int it = 0; it < 5; it++) prf_kernel<<<grid, TPB, smem>>>(key, out, nvec);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  uint4 h[2];
  cudaMemcpy(h, out, 32, cudaMemcpyDeviceToHost);
  printf("abs=%d tpb=%d nb=%d grid=%d: %.2f G blocks/s (%.1f GB/s), first block %08x %08x %08x %08x, err=%s\n", AES_ABS_TABLE, TPB, NB, grid,
         5.0 * nvec / ms / 1e6, 5.0 * nvec * 16 / ms / 1e6, h[0].x, h[0].y, h[0].z, h[0].w, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
int mb_fail(const mb_ctx *, int code, const std::string &msg) { fprintf(stderr, "%s\n", msg.c_str()); return code; }
