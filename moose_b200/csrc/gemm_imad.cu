// Ring GEMM mod 2^64 / 2^128 on the CUDA cores (64-bit IMAD chains).
//
// DotOp::ring_kernel (moose/src/host/ops.rs:1754-1851) for the shapes the int8
// tensor-pipe kernel (gemm_limb.cu) is the wrong tool for:
//  * skinny outputs (N <= 32: LogReg [B,513]x[513,2], the last MLP layer
//    [B,512]x[512,10]) -- these stream A once and are HBM bound;
//  * small / ragged problems (unit tests, KATs, rank-1 operands).
// Integer arithmetic mod 2^k is associative and commutative, so the summation
// order does not affect the result: bit-exact against ndarray's generic dot.
#include "vecops.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16, TPB = 256;

// Tiled kernel: 64x64 output tile per CTA, 4x4 outputs per thread.
template <typename T>
__global__ void __launch_bounds__(TPB) gemm_tiled_kernel(const T *__restrict__ A, const T *__restrict__ B, T *__restrict__ C,
                                                         int64_t M, int64_t K, int64_t N, bool accumulate) {
  __shared__ T sA[BK][BM + 1];
  __shared__ T sB[BK][BN];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;
  T acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0;

  for (int64_t k0 = 0; k0 < K; k0 += BK) {
    // A tile: 64 rows x 16 k  (k contiguous in memory)
    for (int idx = threadIdx.x; idx < BM * BK; idx += TPB) {
      int r = idx / BK, kk = idx % BK;
      int64_t gm = m0 + r, gk = k0 + kk;
      sA[kk][r] = (gm < M && gk < K) ? A[gm * K + gk] : (T)0;
    }
    // B tile: 16 k x 64 cols (n contiguous)
    for (int idx = threadIdx.x; idx < BK * BN; idx += TPB) {
      int kk = idx / BN, c = idx % BN;
      int64_t gk = k0 + kk, gn = n0 + c;
      sB[kk][c] = (gk < K && gn < N) ? B[gk * N + gn] : (T)0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; kk++) {
      T a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = sA[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; j++) b[j] = sB[kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] += a[i] * b[j];
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    int64_t gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int64_t gn = n0 + tx + 16 * j;
      if (gn < N) {
        T v = acc[i][j];
        if (accumulate) v += C[gm * N + gn];
        C[gm * N + gn] = v;
      }
    }
  }
}

// Skinny kernel (N <= 32): one warp per output row, lanes stride over K so the A
// row is read with coalesced 8/16-byte loads exactly once; B (K x N, small) stays
// in L1/L2.  Columns are processed in chunks of NC to bound registers.
template <typename T, int NC>
__global__ void __launch_bounds__(TPB) gemm_skinny_kernel(const T *__restrict__ A, const T *__restrict__ B, T *__restrict__ C,
                                                          int64_t M, int64_t K, int64_t N, bool accumulate) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * TPB + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * TPB) >> 5;
  for (int64_t m = warp; m < M; m += nwarps) {
    const T *a = A + m * K;
    for (int64_t n0 = 0; n0 < N; n0 += NC) {
      T acc[NC];
#pragma unroll
      for (int j = 0; j < NC; j++) acc[j] = 0;
      for (int64_t k = lane; k < K; k += 32) {
        T av = a[k];
#pragma unroll
        for (int j = 0; j < NC; j++)
          if (n0 + j < N) acc[j] += av * B[k * N + n0 + j];
      }
#pragma unroll
      for (int j = 0; j < NC; j++) {
        T v = acc[j];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          uint64_t lo = __shfl_xor_sync(0xffffffffu, (uint64_t)v, off);
          if (sizeof(T) == 16) {
            uint64_t hi = __shfl_xor_sync(0xffffffffu, (uint64_t)((u128)v >> 64), off);
            v += (T)(((u128)hi << 64) | lo);
          } else {
            v += (T)lo;
          }
        }
        if (lane == 0 && n0 + j < N) {
          if (accumulate) v += C[m * N + n0 + j];
          C[m * N + n0 + j] = v;
        }
      }
    }
  }
}

// Replicated dot with a skinny right operand, all three parties in one pass: every share of x is read ONCE
// (the party-by-party form reads each share twice) and the left operand may be a virtual concatenation along K of
// up to four parts (linear predictors: [bias column, features]).  One warp per row:
//   v_i[m, :] = sum_k x_i[m, k] * (y_i + y_{i+1})[k, :] + x_{i+1}[m, k] * y_i[k, :]        (arith.rs:461-469)
template <typename T, int NC>
__global__ void __launch_bounds__(TPB) rep_dot_skinny_kernel(RepDotParts parts, const T *__restrict__ yb0, const T *__restrict__ yb1,
                                                              const T *__restrict__ yb2, const T *__restrict__ y0, const T *__restrict__ y1,
                                                              const T *__restrict__ y2, T *__restrict__ z0, T *__restrict__ z1,
                                                              T *__restrict__ z2, int64_t M, int64_t N) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * TPB + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * TPB) >> 5;
  for (int64_t m = warp; m < M; m += nwarps) {
    for (int64_t n0 = 0; n0 < N; n0 += NC) {
      T acc[3][NC];
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < NC; j++) acc[i][j] = 0;
      for (int p = 0; p < parts.n; p++) {
        const int64_t K = parts.p[p].K, r0 = parts.p[p].krow0;
        const T *a0p = (const T *)parts.p[p].x[0] + m * K, *a1p = (const T *)parts.p[p].x[1] + m * K,
                *a2p = (const T *)parts.p[p].x[2] + m * K;
        for (int64_t k = lane; k < K; k += 32) {
          const T a0 = a0p[k], a1 = a1p[k], a2 = a2p[k];
          const int64_t row = (r0 + k) * N + n0;
#pragma unroll
          for (int j = 0; j < NC; j++)
            if (n0 + j < N) {
              acc[0][j] += a0 * yb0[row + j] + a1 * y0[row + j];
              acc[1][j] += a1 * yb1[row + j] + a2 * y1[row + j];
              acc[2][j] += a2 * yb2[row + j] + a0 * y2[row + j];
            }
        }
      }
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < NC; j++) {
          T v = acc[i][j];
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) {
            uint64_t lo = __shfl_xor_sync(0xffffffffu, (uint64_t)v, off);
            if (sizeof(T) == 16) {
              uint64_t hi = __shfl_xor_sync(0xffffffffu, (uint64_t)((u128)v >> 64), off);
              v += (T)(((u128)hi << 64) | lo);
            } else {
              v += (T)lo;
            }
          }
          if (lane == 0 && n0 + j < N) (i == 0 ? z0 : i == 1 ? z1 : z2)[m * N + n0 + j] = v;
        }
    }
  }
}

// Narrow right operand on Ring64 (N <= 16, even; the last layer of the MLP: [65536, 512] x [512, 10]): ONE LANE PER ROW.
// The warp-per-row kernel above spends one load per product on the weights; here a lane keeps the 3 x N accumulators
// of its row in registers and walks k, so the weights are warp-uniform (broadcast) vector loads shared by 32 rows and
// the x shares are read one cache line per lane per 16 k.  v_i = x_i . (y_i + y_{i+1}) + x_{i+1} . y_i  (arith.rs:461-469)
template <int N>
__global__ void __launch_bounds__(128) rep_dot_rowlane_kernel(const uint64_t *__restrict__ x0, const uint64_t *__restrict__ x1,
                                                              const uint64_t *__restrict__ x2, const uint64_t *__restrict__ yb0,
                                                              const uint64_t *__restrict__ yb1, const uint64_t *__restrict__ yb2,
                                                              const uint64_t *__restrict__ y0, const uint64_t *__restrict__ y1,
                                                              const uint64_t *__restrict__ y2, uint64_t *__restrict__ z0,
                                                              uint64_t *__restrict__ z1, uint64_t *__restrict__ z2, int64_t M, int64_t K) {
  // x tile of the CTA's 128 rows x KC columns per share, staged through shared memory: global loads run along k
  // (coalesced), a lane then reads its own row (row stride KC + 1 words of 8 bytes: conflict free)
  constexpr int KC = 8, LD = KC + 1;   // 3 x 128 x 9 x 8 bytes = 27 KB of static shared memory
  __shared__ uint64_t sx[3][128 * LD];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t mb = (int64_t)blockIdx.x * 128; mb < M; mb += (int64_t)gridDim.x * 128) {
    const int64_t m = mb + threadIdx.x;
    uint64_t a0[N], a1[N], a2[N];
#pragma unroll
    for (int j = 0; j < N; j++) a0[j] = a1[j] = a2[j] = 0;
    for (int64_t k0 = 0; k0 < K; k0 += KC) {
      __syncthreads();
      // a 32-lane access loads four rows of KC = 8 columns (64-byte segments)
      for (int r = warp * 4 + (lane >> 3); r < 128; r += 16) {
        const int64_t gm = mb + r, gk = k0 + (lane & 7);
        const bool ok = gm < M && gk < K;
        sx[0][r * LD + (lane & 7)] = ok ? x0[gm * K + gk] : 0;
        sx[1][r * LD + (lane & 7)] = ok ? x1[gm * K + gk] : 0;
        sx[2][r * LD + (lane & 7)] = ok ? x2[gm * K + gk] : 0;
      }
      __syncthreads();
      const int kc = (int)(K - k0 < KC ? K - k0 : KC);
      for (int kk = 0; kk < kc; kk++) {
        const int64_t k = k0 + kk;
        const uint64_t u0 = sx[0][threadIdx.x * LD + kk], u1 = sx[1][threadIdx.x * LD + kk], u2 = sx[2][threadIdx.x * LD + kk];
        const ulonglong2 *rb0 = (const ulonglong2 *)(yb0 + k * N), *rb1 = (const ulonglong2 *)(yb1 + k * N), *rb2 = (const ulonglong2 *)(yb2 + k * N);
        const ulonglong2 *r0 = (const ulonglong2 *)(y0 + k * N), *r1 = (const ulonglong2 *)(y1 + k * N), *r2 = (const ulonglong2 *)(y2 + k * N);
#pragma unroll
        for (int j = 0; j < N / 2; j++) {
          const ulonglong2 b0 = __ldg(rb0 + j), b1 = __ldg(rb1 + j), b2 = __ldg(rb2 + j);
          const ulonglong2 c0 = __ldg(r0 + j), c1 = __ldg(r1 + j), c2 = __ldg(r2 + j);
          a0[2 * j] += u0 * b0.x + u1 * c0.x;  a0[2 * j + 1] += u0 * b0.y + u1 * c0.y;
          a1[2 * j] += u1 * b1.x + u2 * c1.x;  a1[2 * j + 1] += u1 * b1.y + u2 * c1.y;
          a2[2 * j] += u2 * b2.x + u0 * c2.x;  a2[2 * j + 1] += u2 * b2.y + u0 * c2.y;
        }
      }
    }
    if (m < M) {
#pragma unroll
      for (int j = 0; j < N; j++) {
        z0[m * N + j] = a0[j];
        z1[m * N + j] = a1[j];
        z2[m * N + j] = a2[j];
      }
    }
  }
}

// measured on the LogReg input (10^6 x 512 Ring128): 1024 threads / no unroll 17.0 ms, 512 / 2 15.6 ms, 512 / 4 16.2 ms,
// 768 / 2 17.9 ms (spills)
#ifndef SKINNY_PRF_TPB
#define SKINNY_PRF_TPB 512
#endif
#ifndef SKINNY_PRF_UNROLL
#define SKINNY_PRF_UNROLL 2
#endif
// One deferred part of one row (owner O, compile time): party O holds (r, d), party O+1 holds (d, 0), party O+2 holds
// (0, r) with r = PRF block of the element, d = plain - r -- four products per element and output column:
//   v_O     += r * (y_O + y_{O+1}) + d * y_O,   v_{O+1} += d * (y_{O+1} + y_{O+2}),   v_{O+2} += r * y_{O+2}
template <typename T, int NC, int O>
__device__ __forceinline__ void deferred_part(const AesLane &L, const AesKey &key, const T *__restrict__ xp, int64_t e0, int64_t K, int64_t r0,
                                              int64_t N, int lane, const T *__restrict__ ybo, const T *__restrict__ ybn,
                                              const T *__restrict__ yo, const T *__restrict__ yp, T (&acc)[3][NC]) {
  constexpr int EPV = 16 / (int)sizeof(T);
  constexpr int VK = sizeof(T) == 16 ? VK_RING128 : VK_RING64;
  constexpr int UNROLL = SKINNY_PRF_UNROLL;
#pragma unroll UNROLL
  for (int64_t k = (int64_t)lane * EPV; k < K; k += 32 * EPV) {
    const vec_t rv = prf_vec<VK>(L, key, (uint64_t)((e0 + k) / EPV));
#pragma unroll
    for (int e = 0; e < EPV; e++) {
      T r;
      if (sizeof(T) == 16) r = (T)v128(rv);
      else r = (T)(e == 0 ? rv.x : rv.y);
      const T d = xp[k + e] - r;
      const int64_t row = (r0 + k + e) * N;
#pragma unroll
      for (int j = 0; j < NC; j++)
        if (j < N) {
          acc[O][j] += r * ybo[row + j] + d * yo[row + j];
          acc[(O + 1) % 3][j] += d * ybn[row + j];
          acc[(O + 2) % 3][j] += r * yp[row + j];
        }
    }
  }
}

// The same with the PRF in the kernel: parts with lazy.owner[p] >= 0 are read from their PLAIN tensor and shared on the
// fly (x_owner = PRF block of the element's vector, x_{owner+1} = plain - x_owner, x_{owner+2} = 0), exactly the values
// share_kernel (rep.cu) would have written -- 8 GB per share at BASELINE config 3 that are neither written nor read.
// One CTA per SM (the AES table), one warp per row; a lane takes the EPV = 16 / sizeof(T) elements of one AES block.
template <typename T, int NC>
__global__ void __launch_bounds__(SKINNY_PRF_TPB) rep_dot_skinny_prf_kernel(RepDotParts parts, const __grid_constant__ RepDotDeferred lazy,
                                                                           const T *__restrict__ yb0, const T *__restrict__ yb1,
                                                                           const T *__restrict__ yb2, const T *__restrict__ y0,
                                                                           const T *__restrict__ y1, const T *__restrict__ y2,
                                                                           T *__restrict__ z0, T *__restrict__ z1, T *__restrict__ z2,
                                                                           int64_t M, int64_t N) {
  extern __shared__ __align__(16) uint32_t sT[];
  aes_load_table(sT);
  const AesLane L = aes_lane(sT);
  constexpr int EPV = 16 / (int)sizeof(T);
  constexpr int VK = sizeof(T) == 16 ? VK_RING128 : VK_RING64;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t m = warp; m < M; m += nwarps) {
    T acc[3][NC];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < NC; j++) acc[i][j] = 0;
    for (int p = 0; p < parts.n; p++) {
      const int64_t K = parts.p[p].K, r0 = parts.p[p].krow0;
      const int o = lazy.owner[p];
      if (o < 0) {
        const T *a0p = (const T *)parts.p[p].x[0] + m * K, *a1p = (const T *)parts.p[p].x[1] + m * K,
                *a2p = (const T *)parts.p[p].x[2] + m * K;
        for (int64_t k = lane; k < K; k += 32) {
          const T a0 = a0p[k], a1 = a1p[k], a2 = a2p[k];
          const int64_t row = (r0 + k) * N;
#pragma unroll
          for (int j = 0; j < NC; j++)
            if (j < N) {
              acc[0][j] += a0 * yb0[row + j] + a1 * y0[row + j];
              acc[1][j] += a1 * yb1[row + j] + a2 * y1[row + j];
              acc[2][j] += a2 * yb2[row + j] + a0 * y2[row + j];
            }
        }
      } else {
        const T *xp = (const T *)lazy.plain[p] + m * K;
        if (o == 0) deferred_part<T, NC, 0>(L, lazy.key[p], xp, m * K, K, r0, N, lane, yb0, yb1, y0, y2, acc);
        else if (o == 1) deferred_part<T, NC, 1>(L, lazy.key[p], xp, m * K, K, r0, N, lane, yb1, yb2, y1, y0, acc);
        else deferred_part<T, NC, 2>(L, lazy.key[p], xp, m * K, K, r0, N, lane, yb2, yb0, y2, y1, acc);
      }
    }
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < NC; j++) {
        T v = acc[i][j];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          uint64_t lo = __shfl_xor_sync(0xffffffffu, (uint64_t)v, off);
          if (sizeof(T) == 16) {
            uint64_t hi = __shfl_xor_sync(0xffffffffu, (uint64_t)((u128)v >> 64), off);
            v += (T)(((u128)hi << 64) | lo);
          } else {
            v += (T)lo;
          }
        }
        if (lane == 0 && j < N) (i == 0 ? z0 : i == 1 ? z1 : z2)[m * N + j] = v;
      }
  }
}

}  // namespace

bool k_rep_dot_skinny_supported(int64_t M, int64_t N) { return N <= 32 && M >= 64; }

// the PRF is evaluated once per element only if all output columns fit one register chunk; rows must start on AES blocks
bool k_rep_dot_skinny_deferred_supported(mb_dtype dtype, const RepDotParts &parts, int64_t N) {
  const int epv = dtype == MB_RING128 ? 1 : 2, nc = dtype == MB_RING128 ? 2 : 4;
  if (N > nc) return false;
  for (int p = 0; p < parts.n; p++)
    if (parts.p[p].K % epv) return false;
  return true;
}

int k_rep_dot_skinny_deferred(mb_ctx *ctx, mb_dtype dtype, const RepDotParts &parts, const RepDotDeferred &lazy, const void *const ysum[3],
                              const void *const y[3], void *const z[3], int64_t M, int64_t N) {
  if (M == 0 || N == 0) return MB_OK;
  const mb_aes_cfg_t cfg = mb_aes_cfg(ctx, M * 32, SKINNY_PRF_TPB);   // one warp per row
  double kt = 0, kprf = 0, kmat = 0;
  for (int i = 0; i < parts.n; i++) {
    kt += (double)parts.p[i].K;
    if (lazy.owner[i] >= 0) kprf += (double)parts.p[i].K; else kmat += (double)parts.p[i].K;
  }
  const double w = dtype == MB_RING128 ? 16.0 : 8.0;
  if (dtype == MB_RING128) {
    MB_CUDA(ctx, cudaFuncSetAttribute(rep_dot_skinny_prf_kernel<u128, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, AES_SMEM_BYTES));
    rep_dot_skinny_prf_kernel<u128, 2><<<cfg.grid_dim, cfg.block_dim, AES_SMEM_BYTES, ctx->stream>>>(
        parts, lazy, (const u128 *)ysum[0], (const u128 *)ysum[1], (const u128 *)ysum[2], (const u128 *)y[0], (const u128 *)y[1],
        (const u128 *)y[2], (u128 *)z[0], (u128 *)z[1], (u128 *)z[2], M, N);
  } else {
    MB_CUDA(ctx, cudaFuncSetAttribute(rep_dot_skinny_prf_kernel<uint64_t, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, AES_SMEM_BYTES));
    rep_dot_skinny_prf_kernel<uint64_t, 4><<<cfg.grid_dim, cfg.block_dim, AES_SMEM_BYTES, ctx->stream>>>(
        parts, lazy, (const uint64_t *)ysum[0], (const uint64_t *)ysum[1], (const uint64_t *)ysum[2], (const uint64_t *)y[0],
        (const uint64_t *)y[1], (const uint64_t *)y[2], (uint64_t *)z[0], (uint64_t *)z[1], (uint64_t *)z[2], M, N);
  }
  // plain operand read once, materialised parts three shares once, result written once; one AES block per 16 plain bytes
  MB_LAUNCH_CHECK_W(ctx, "rep_dot_skinny_deferred", w * ((double)M * (kprf + 3.0 * kmat) + 3.0 * M * N), (double)M * kprf * w / 16.0, 0);
  return MB_OK;
}

int k_rep_dot_skinny(mb_ctx *ctx, mb_dtype dtype, const RepDotParts &parts, const void *const ysum[3], const void *const y[3],
                     void *const z[3], int64_t M, int64_t N) {
  if (M == 0 || N == 0) return MB_OK;
  if (dtype == MB_RING64 && parts.n == 1 && M >= 128 * 2 * (int64_t)ctx->sm_count && N % 2 == 0 && N >= 2 && N <= 16) {   // one lane per row: needs >= 2 CTAs of 128 rows per SM
    const uint64_t *x0 = (const uint64_t *)parts.p[0].x[0], *x1 = (const uint64_t *)parts.p[0].x[1], *x2 = (const uint64_t *)parts.p[0].x[2];
    const int64_t K = parts.p[0].K;
    const int g = (int)std::min<int64_t>((M + 127) / 128, (int64_t)ctx->sm_count * 8);
#define ROWLANE(NN)                                                                                                                  \
  case NN:                                                                                                                           \
    rep_dot_rowlane_kernel<NN><<<g, 128, 0, ctx->stream>>>(x0, x1, x2, (const uint64_t *)ysum[0], (const uint64_t *)ysum[1],           \
                                                           (const uint64_t *)ysum[2], (const uint64_t *)y[0], (const uint64_t *)y[1], \
                                                           (const uint64_t *)y[2], (uint64_t *)z[0], (uint64_t *)z[1], (uint64_t *)z[2], M, K); \
    break;
    switch ((int)N) {
      ROWLANE(2) ROWLANE(4) ROWLANE(6) ROWLANE(8) ROWLANE(10) ROWLANE(12) ROWLANE(14) ROWLANE(16)
    }
#undef ROWLANE
    MB_LAUNCH_CHECK_W(ctx, "rep_dot_rowlane", 3.0 * 8.0 * ((double)M * K + (double)M * N), 0, 0);
    return MB_OK;
  }
  const int g = (int)std::min<int64_t>((M * 32 + TPB - 1) / TPB, (int64_t)ctx->sm_count * 16);
  if (dtype == MB_RING128)
    rep_dot_skinny_kernel<u128, 2><<<g, TPB, 0, ctx->stream>>>(parts, (const u128 *)ysum[0], (const u128 *)ysum[1], (const u128 *)ysum[2],
                                                              (const u128 *)y[0], (const u128 *)y[1], (const u128 *)y[2], (u128 *)z[0],
                                                              (u128 *)z[1], (u128 *)z[2], M, N);
  else
    rep_dot_skinny_kernel<uint64_t, 4><<<g, TPB, 0, ctx->stream>>>(parts, (const uint64_t *)ysum[0], (const uint64_t *)ysum[1],
                                                                  (const uint64_t *)ysum[2], (const uint64_t *)y[0], (const uint64_t *)y[1],
                                                                  (const uint64_t *)y[2], (uint64_t *)z[0], (uint64_t *)z[1],
                                                                  (uint64_t *)z[2], M, N);
  {
    const double w = dtype == MB_RING128 ? 16.0 : 8.0;
    double kt = 0;
    for (int i = 0; i < parts.n; i++) kt += (double)parts.p[i].K;
    // every share of the left operand read once, the result written once (the weights stay in L2)
    MB_LAUNCH_CHECK_W(ctx, "rep_dot_skinny", 3.0 * w * ((double)M * kt + (double)M * N), 0, 0);
  }
  return MB_OK;
}

int k_gemm_imad(mb_ctx *ctx, mb_dtype dtype, const void *A, const void *B, void *C, int64_t M, int64_t K, int64_t N,
                bool accumulate) {
  if (M == 0 || N == 0) return MB_OK;
  if (K == 0) {
    if (!accumulate) MB_CUDA(ctx, cudaMemsetAsync(C, 0, (size_t)(M * N) * (dtype == MB_RING128 ? 16 : 8), ctx->stream));
    return MB_OK;
  }
  if (N <= 32 && M >= 64) {
    int64_t warps = M;
    int g = (int)std::min<int64_t>((warps * 32 + TPB - 1) / TPB, (int64_t)ctx->sm_count * 16);
    if (dtype == MB_RING128)
      gemm_skinny_kernel<u128, 2><<<g, TPB, 0, ctx->stream>>>((const u128 *)A, (const u128 *)B, (u128 *)C, M, K, N, accumulate);
    else
      gemm_skinny_kernel<uint64_t, 4><<<g, TPB, 0, ctx->stream>>>((const uint64_t *)A, (const uint64_t *)B, (uint64_t *)C, M, K, N, accumulate);
    MB_LAUNCH_CHECK_W(ctx, "gemm_skinny", (dtype == MB_RING128 ? 16.0 : 8.0) * ((double)M * K + (double)K * N + (double)M * N), 0, 0);
    return MB_OK;
  }
  dim3 grid((unsigned)((N + BN - 1) / BN), (unsigned)((M + BM - 1) / BM));
  if (dtype == MB_RING128)
    gemm_tiled_kernel<u128><<<grid, TPB, 0, ctx->stream>>>((const u128 *)A, (const u128 *)B, (u128 *)C, M, K, N, accumulate);
  else
    gemm_tiled_kernel<uint64_t><<<grid, TPB, 0, ctx->stream>>>((const uint64_t *)A, (const uint64_t *)B, (uint64_t *)C, M, K, N, accumulate);
  MB_LAUNCH_CHECK_W(ctx, "gemm_tiled", (dtype == MB_RING128 ? 16.0 : 8.0) * ((double)M * K + (double)K * N + (double)M * N), 0, 0);
  return MB_OK;
}
