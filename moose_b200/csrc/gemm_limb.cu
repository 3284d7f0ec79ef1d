// Ring GEMM mod 2^64 / 2^128 on the Blackwell int8 tensor pipe (tcgen05.mma kind::i8).
//
// DotOp::ring_kernel (moose/src/host/ops.rs:1754-1851) for large dense shapes.  A ring
// element is split into L = 8 (Ring64) or 16 (Ring128) unsigned byte limbs,
//     a = sum_i a_i 2^(8i),  b = sum_j b_j 2^(8j),
//     a*b mod 2^k = sum_{d<L} 2^(8d) * C_d mod 2^k,   C_d = sum_{i+j=d} a_i*b_j
// so the contraction becomes L(L+1)/2 u8 x u8 -> s32 limb GEMMs (36 / 136), grouped by
// the diagonal d = i+j they accumulate into.  One CTA owns a 128 x BLOCK_N output tile
// and walks the diagonals in PAIRS (d, d+1): both accumulators live in TMEM (2 x 256
// columns) and the limb planes are visited in zig-zag order
//     (A_0,B_{d+1}) (A_0,B_d) (A_1,B_d) (A_1,B_{d-1}) ... (A_{d+1},B_0)
// so every operand tile TMA brings into shared memory feeds two MMAs: half the L2->SM
// traffic of the naive order (the 8-bit MMA rate makes operand delivery, not the tensor
// pipe, the first bound).
//
// Exactness: u8*u8 <= 65025 and the tensor core accumulates mod 2^32 (no saturation), so
// an accumulator is exact while it holds <= 66051 products.  Diagonals whose shift 8d
// leaves <= 32 significant bits (8d >= k-32) may wrap freely; the others are cut into
// passes of bounded length by the host-side planner.  Each pass writes one u64 partial
// C_d + (C_{d+1} << 8) per output element; a second kernel recombines
//     C = sum_p partial_p << (8 d_p)   mod 2^k
// (HBM traffic of the partials is a few % of the MMA time at K = 4096).
//
// Scheduling: the unit of work is one (batch, tile, pass) triple -- passes of a tile are independent because
// each writes its own partial slot.  Passes are sorted by cost (MMA steps) and the flat work list is dealt to
// the persistent CTAs in snake order (0..G-1, G-1..0, ...), a static longest-first schedule that removes the
// wave tail of a tile-granular launch; the three parties of a replicated dot are one batched launch.
//
// Warp roles (192 threads): warps 0-3 epilogue (TMEM -> registers -> global), warp 4 TMA
// producer, warp 5 TMEM allocator + single-thread MMA issuer.
#include <cuda.h>

#include "common.h"

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 128;  // bytes of K per shared-memory tile row = one 128B swizzle atom
constexpr int UMMA_K = 32;    // u8 elements per tcgen05.mma
constexpr int MAX_BLOCK_N = 256;
constexpr int A_SLOTS = 4, B_SLOTS = 4;
constexpr int A_TILE_BYTES = BLOCK_M * BLOCK_K;      // 16 KB
constexpr int B_TILE_BYTES = MAX_BLOCK_N * BLOCK_K;  // 32 KB
constexpr int SMEM_BYTES = A_SLOTS * A_TILE_BYTES + B_SLOTS * B_TILE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int NUM_THREADS = 192;
constexpr uint32_t ACC_LIMIT = 66051;  // floor(2^32 / 255^2)

constexpr int MAX_PASSES = 64, MAX_UNITS = 128;

struct PassDesc {
  short d;          // lower diagonal of the pair (d, d+1); d+1 may be >= L (then only d is live)
  short unit_begin; // [unit_begin, unit_end) into the unit list
  short unit_end;
  short pad;
};
struct UnitDesc {
  int term;      // operand pair
  int kb_begin;  // k-block range
  int kb_end;
};

constexpr int MAX_BATCH = 4;

struct GemmParams {
  PassDesc passes[MAX_PASSES];   // sorted by decreasing cost; index = partial slot
  UnitDesc units[MAX_UNITS];
  short plane_a[MAX_BATCH][2];   // first limb plane of the A operand of (batch, term)
  short plane_b[MAX_BATCH][2];
  int num_passes;
  int nbatch;
  int L;            // limbs per element
  int M, N;
  int block_n;
  int tiles_m, tiles_n;
  int total_items;  // num_passes * nbatch * tiles_m * tiles_n
  unsigned long long *partials;  // [nbatch][num_passes][M][N] u64
  int *round_barrier;  // 2-CTA kernel: one zeroed counter per scheduling round (nullptr = free running)
  int barrier_rounds;  // number of rounds; slot [barrier_rounds] is the `broken` flag (bounded waits)
};

// Work item j of CTA c in snake order over the flat list [pass rank][batch][tile].
struct WorkItem {
  int pass, batch, m0, n0;
};
__device__ __forceinline__ bool next_item(const GemmParams &p, int j, WorkItem &w) {
  const int G = (int)gridDim.x;
  const int c = (j & 1) ? G - 1 - (int)blockIdx.x : (int)blockIdx.x;
  const long long idx = (long long)j * G + c;
  if (idx >= p.total_items) return false;
  const int tiles = p.tiles_m * p.tiles_n;
  const int per_pass = p.nbatch * tiles;
  w.pass = (int)(idx / per_pass);
  const int rem = (int)(idx % per_pass);
  w.batch = rem / tiles;
  const int tile = rem % tiles;
  w.m0 = (tile / p.tiles_n) * BLOCK_M;
  w.n0 = (tile % p.tiles_n) * p.block_n;
  return true;
}

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  uint32_t spins = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!done && ++spins > (1u << 26)) __trap();  // watchdog: turn a protocol bug into an error, not a hang
  }
}
// Same, for the 128 epilogue threads: they wait for whole passes (10^5..10^6 cycles), so they sleep between
// polls instead of competing with the TMA / MMA issuing threads for issue slots.
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  uint32_t spins = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!done) {
      __nanosleep(256);
      if (++spins > (1u << 24)) __trap();
    }
  }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem], u8 x u8 -> s32, M = 128, N from idesc, K = 32.
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrive once all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128B-swizzled operand tile (rows x 128 bytes, 8-row groups 1024 bytes apart):
// start address >> 4 | LBO (ignored for swizzled K-major) | SBO = 1024 >> 4 | version 1 | SWIZZLE_128B.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
  uint64_t desc = 0;
  desc |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  desc |= (uint64_t)1 << 16;
  desc |= (uint64_t)(1024 >> 4) << 32;
  desc |= (uint64_t)1 << 46;
  desc |= (uint64_t)2 << 61;
  return desc;
}

// ---------------------------------------------------------------- main kernel
__global__ void __launch_bounds__(NUM_THREADS, 1)
limb_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the 128B swizzle atoms
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *sA = smem;
  uint8_t *sB = smem + A_SLOTS * A_TILE_BYTES;
  uint64_t *bars = (uint64_t *)(sB + B_SLOTS * B_TILE_BYTES);
  // barrier layout: fullA[4] emptyA[4] fullB[4] emptyB[4] tmem_full tmem_empty
  const uint32_t bar_base = smem_u32(bars);
  auto fullA = [&](int s) { return bar_base + 8u * s; };
  auto emptyA = [&](int s) { return bar_base + 8u * (A_SLOTS + s); };
  auto fullB = [&](int s) { return bar_base + 8u * (2 * A_SLOTS + s); };
  auto emptyB = [&](int s) { return bar_base + 8u * (2 * A_SLOTS + B_SLOTS + s); };
  const uint32_t tmem_full = bar_base + 8u * (2 * A_SLOTS + 2 * B_SLOTS);
  const uint32_t tmem_empty = tmem_full + 8u;
  uint32_t *tmem_slot = (uint32_t *)(bars + 2 * A_SLOTS + 2 * B_SLOTS + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < A_SLOTS; s++) { mbar_init(fullA(s), 1); mbar_init(emptyA(s), 1); }
    for (int s = 0; s < B_SLOTS; s++) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), 1); }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 128);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const uint32_t b_tile_bytes = (uint32_t)p.block_n * BLOCK_K;
  const int L = p.L;

  if (warp == 4) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int a_it = 0, b_it = 0;  // running slot counters
      WorkItem w;
      for (int item = 0; next_item(p, item, w); item++) {
        const PassDesc pd = p.passes[w.pass];
        const int d = pd.d;
        const bool two = d + 1 < L;              // pair has an upper diagonal
        const int last = two ? 2 * d + 2 : -1;   // last zig-zag step (two diagonals)
        for (int u = pd.unit_begin; u < pd.unit_end; u++) {
          const UnitDesc ud = p.units[u];
          const int pa0 = p.plane_a[w.batch][ud.term], pb0 = p.plane_b[w.batch][ud.term];
          for (int kb = ud.kb_begin; kb < ud.kb_end; kb++) {
            const int k0 = kb * BLOCK_K;
            if (two) {
              for (int s = 0; s <= last; s++) {
                const int t = s >> 1;
                if ((s & 1) == 0) {  // new A_t at even steps
                  const int slot = a_it % A_SLOTS;
                  mbar_wait(emptyA(slot), ((a_it / A_SLOTS) & 1) ^ 1);
                  mbar_expect_tx(fullA(slot), A_TILE_BYTES);
                  tma_load_3d(smem_u32(sA + slot * A_TILE_BYTES), &map_a, fullA(slot), k0, w.m0, pa0 + t);
                  a_it++;
                }
                if (s == 0 || (s & 1) == 1) {  // new B at step 0 (B_{d+1}) and odd steps (B_{d-t})
                  const int j = (s == 0) ? d + 1 : d - t;
                  const int slot = b_it % B_SLOTS;
                  mbar_wait(emptyB(slot), ((b_it / B_SLOTS) & 1) ^ 1);
                  mbar_expect_tx(fullB(slot), b_tile_bytes);
                  tma_load_3d(smem_u32(sB + slot * B_TILE_BYTES), &map_b, fullB(slot), k0, w.n0, pb0 + j);
                  b_it++;
                }
              }
            } else {
              // single (last, odd-L) diagonal: plain pairs (i, d-i)
              for (int i = 0; i <= d; i++) {
                int slot = a_it % A_SLOTS;
                mbar_wait(emptyA(slot), ((a_it / A_SLOTS) & 1) ^ 1);
                mbar_expect_tx(fullA(slot), A_TILE_BYTES);
                tma_load_3d(smem_u32(sA + slot * A_TILE_BYTES), &map_a, fullA(slot), k0, w.m0, pa0 + i);
                a_it++;
                slot = b_it % B_SLOTS;
                mbar_wait(emptyB(slot), ((b_it / B_SLOTS) & 1) ^ 1);
                mbar_expect_tx(fullB(slot), b_tile_bytes);
                tma_load_3d(smem_u32(sB + slot * B_TILE_BYTES), &map_b, fullB(slot), k0, w.n0, pb0 + d - i);
                b_it++;
              }
            }
          }
        }
      }
    }
  } else if (warp == 5) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      // instruction descriptor: D = S32, A = B = U8, both K-major, N = block_n, M = 128
      const uint32_t idesc = (2u << 4) | ((uint32_t)(p.block_n >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
      int a_it = 0, b_it = 0;
      int a_slot = 0, b_slot = 0;
      uint32_t acc_phase = 0;
      WorkItem w;
      for (int item = 0; next_item(p, item, w); item++) {
        const PassDesc pd = p.passes[w.pass];
        const int d = pd.d;
        const bool two = d + 1 < L;
        const int last = 2 * d + 2;
        // wait until the epilogue has drained the accumulators of the previous work item
        mbar_wait(tmem_empty, acc_phase ^ 1);
        tcgen05_fence_after();
        bool first_lo = true, first_hi = true;  // first MMA into acc(d) / acc(d+1) of this pass
        for (int u = pd.unit_begin; u < pd.unit_end; u++) {
          const UnitDesc ud = p.units[u];
          for (int kb = ud.kb_begin; kb < ud.kb_end; kb++) {
            const int nsteps = two ? last + 1 : d + 1;
            for (int s = 0; s < nsteps; s++) {
              bool new_a, new_b, hi, rel_a, rel_b;
              if (two) {
                new_a = (s & 1) == 0;
                new_b = s == 0 || (s & 1) == 1;
                hi = (s & 1) == 0;             // even steps accumulate into diagonal d+1
                rel_a = (s & 1) == 1 || s == last;
                rel_b = (s & 1) == 0;
              } else {
                new_a = new_b = rel_a = rel_b = true;
                hi = false;
              }
              if (new_a) {
                a_slot = a_it % A_SLOTS;
                mbar_wait(fullA(a_slot), (a_it / A_SLOTS) & 1);
                a_it++;
              }
              if (new_b) {
                b_slot = b_it % B_SLOTS;
                mbar_wait(fullB(b_slot), (b_it / B_SLOTS) & 1);
                b_it++;
              }
              tcgen05_fence_after();
              const uint64_t adesc = make_smem_desc(smem_u32(sA + a_slot * A_TILE_BYTES));
              const uint64_t bdesc = make_smem_desc(smem_u32(sB + b_slot * B_TILE_BYTES));
              const uint32_t tmem_d = tmem_base + (hi ? (uint32_t)MAX_BLOCK_N : 0u);
              bool &first = hi ? first_hi : first_lo;
#pragma unroll
              for (int k = 0; k < BLOCK_K / UMMA_K; k++) {
                // advance both descriptors by k * 32 bytes inside the swizzle atom
                umma_i8(tmem_d, adesc + (uint64_t)(k * UMMA_K >> 4), bdesc + (uint64_t)(k * UMMA_K >> 4), idesc,
                        (first && k == 0) ? 0u : 1u);
              }
              first = false;
              if (rel_a) umma_commit(emptyA(a_slot));
              if (rel_b) umma_commit(emptyB(b_slot));
            }
          }
        }
        umma_commit(tmem_full);  // accumulators of this work item are complete
        acc_phase ^= 1;
      }
    }
  } else {
    // ===================== epilogue (warps 0-3; warp w owns TMEM lanes 32w..32w+31) =====================
    uint32_t acc_phase = 0;
    WorkItem w;
    for (int item = 0; next_item(p, item, w); item++) {
      const int row = w.m0 + warp * 32 + lane;
      const int n0 = w.n0;
      const bool two = p.passes[w.pass].d + 1 < L;
      mbar_wait_relaxed(tmem_full, acc_phase);
      tcgen05_fence_after();
      unsigned long long *out =
          p.partials + (((size_t)w.batch * p.num_passes + (size_t)w.pass) * p.M + (size_t)row) * p.N + n0;
      for (int c = 0; c < p.block_n; c += 32) {
        uint32_t lo[32], hi[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c;
        tmem_ld32(taddr, lo);
        if (two) tmem_ld32(taddr + MAX_BLOCK_N, hi);
        tmem_ld_wait();
        if (row < p.M) {
          if (n0 + c + 32 <= p.N) {
#pragma unroll
            for (int q = 0; q < 32; q += 2) {
              unsigned long long v0 = (unsigned long long)lo[q] + (two ? ((unsigned long long)hi[q] << 8) : 0ull);
              unsigned long long v1 = (unsigned long long)lo[q + 1] + (two ? ((unsigned long long)hi[q + 1] << 8) : 0ull);
              *(ulonglong2 *)(out + c + q) = make_ulonglong2(v0, v1);
            }
          } else {
#pragma unroll
            for (int q = 0; q < 32; q++) {
              if (n0 + c + q < p.N)
                out[c + q] = (unsigned long long)lo[q] + (two ? ((unsigned long long)hi[q] << 8) : 0ull);
            }
          }
        }
      }
      tcgen05_fence_before();
      mbar_arrive(tmem_empty);
      acc_phase ^= 1;
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 5) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// ---------------------------------------------------------------- 2-CTA kernel (cta_group::2)
// Same algorithm on CTA pairs: a cluster of two CTAs (one TPC) owns a 256 x BLOCK_N tile.  CTA r loads A rows
// [m0 + 128 r, +128) and B rows (n) [n0 + BLOCK_N/2 * r, +BLOCK_N/2) into ITS shared memory; the leader (rank 0)
// issues tcgen05.mma.cta_group::2 with M = 256, which reads A from both CTAs and the two halves of B, and
// accumulates 128 rows in each CTA's TMEM.  Per MMA step each SM now pulls 16 KB from L2 instead of 24 KB
// (B is shared by the pair): the 1-CTA kernel is bound by L2 -> SM operand delivery, not by the tensor pipe.
//   full barriers   : leader's; both CTAs' TMA transfers complete_tx on them (peer bit cleared in the address)
//   empty barriers  : per CTA; released by tcgen05.commit.cta_group::2 multicast to both CTAs
//   tmem_full       : per CTA (multicast commit); tmem_empty: leader's, all 256 epilogue threads arrive remotely
constexpr int A2_SLOTS = 6, B2_SLOTS = 6;
constexpr int B2_TILE_BYTES = (MAX_BLOCK_N / 2) * BLOCK_K;  // 16 KB: half of the B tile per CTA
// epilogue staging: each of the 4 epilogue warps transposes 32 rows x 32 u64 through shared memory (row stride 33)
constexpr int STAGE_STRIDE = 33;
constexpr int STAGE_BYTES = 4 * 32 * STAGE_STRIDE * 8;
constexpr int SMEM2_BYTES = A2_SLOTS * A_TILE_BYTES + B2_SLOTS * B2_TILE_BYTES + 1024 + 256 + STAGE_BYTES;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA into THIS CTA's shared memory, transaction bytes reported to the LEADER CTA's barrier (`leader_bar` is the
// shared::cluster address of that barrier, see leader_address()).
__device__ __forceinline__ uint32_t leader_address(uint32_t local) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(r) : "r"(local));
  return r;
}
__device__ __forceinline__ void tma_load_3d_2sm(uint32_t dst, const CUtensorMap *map, uint32_t leader_bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void umma_i8_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs of the pair once the MMAs issued so far are done
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}
// arrive on the LEADER CTA's copy of a barrier
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, 0;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
      ::"r"(bar)
      : "memory");
}

__device__ __forceinline__ bool next_item_2cta(const GemmParams &p, int j, WorkItem &w) {
  const int G = (int)(gridDim.x >> 1);           // clusters
  const int me = (int)(blockIdx.x >> 1);
  const int c = (j & 1) ? G - 1 - me : me;
  const long long idx = (long long)j * G + c;
  if (idx >= p.total_items) return false;
  const int tiles = p.tiles_m * p.tiles_n;       // tiles of 256 x block_n
  const int per_pass = p.nbatch * tiles;
  w.pass = (int)(idx / per_pass);
  const int rem = (int)(idx % per_pass);
  w.batch = rem / tiles;
  const int tile = rem % tiles;
  w.m0 = (tile / p.tiles_n) * (2 * BLOCK_M);
  w.n0 = (tile % p.tiles_n) * p.block_n;
  return true;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
limb_gemm_kernel_2cta(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                      const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *sA = smem;
  uint8_t *sB = smem + A2_SLOTS * A_TILE_BYTES;
  uint64_t *bars = (uint64_t *)(sB + B2_SLOTS * B2_TILE_BYTES);
  const uint32_t bar_base = smem_u32(bars);
  auto fullA = [&](int s) { return bar_base + 8u * s; };
  auto emptyA = [&](int s) { return bar_base + 8u * (A2_SLOTS + s); };
  auto fullB = [&](int s) { return bar_base + 8u * (2 * A2_SLOTS + s); };
  auto emptyB = [&](int s) { return bar_base + 8u * (2 * A2_SLOTS + B2_SLOTS + s); };
  const uint32_t tmem_full = bar_base + 8u * (2 * A2_SLOTS + 2 * B2_SLOTS);
  const uint32_t tmem_empty = tmem_full + 8u;
  uint32_t *tmem_slot = (uint32_t *)(bars + 2 * A2_SLOTS + 2 * B2_SLOTS + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < A2_SLOTS; s++) { mbar_init(fullA(s), 1); mbar_init(emptyA(s), 1); }
    for (int s = 0; s < B2_SLOTS; s++) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), 1); }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 256);  // 128 epilogue threads of each CTA
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  cluster_sync_all();  // both CTAs' barriers are initialised before anything is signalled across the pair
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int half_n = p.block_n >> 1;
  const uint32_t b_half_bytes = (uint32_t)half_n * BLOCK_K;
  const int L = p.L;

  if (warp == 4) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int a_it = 0, b_it = 0;
      WorkItem w;
      for (int item = 0; next_item_2cta(p, item, w); item++) {
        if (p.round_barrier) {
          // Round barrier: no cluster starts loading work item `item` before every cluster has reached it.
          // Clusters working on tiles that share operand planes then stay within a few MMA steps of each other,
          // so a tile fetched by one is still in L2 when its neighbours ask for it (free-running clusters drift
          // apart by more than the L2 holds and re-read the planes from HBM ~50x).  Only the TMA threads wait.
          // The grid is sized to what cudaOccupancyMaxActiveClusters reports, but nothing GUARANTEES that every
          // cluster is resident (another context or process may hold SMs), so the wait is bounded: a cluster that
          // waits longer than a round can plausibly last raises the `broken` flag (slot [rounds]) and from then on
          // every cluster runs free -- slower through L2, never stuck, and the clusters not yet scheduled get their
          // SMs when the resident ones finish.
          const int G = (int)(gridDim.x >> 1);
          const long long left = (long long)p.total_items - (long long)item * G;
          const int expect = (int)(left < G ? left : G);
          volatile int *bar = p.round_barrier + item;
          volatile int *broken = p.round_barrier + p.barrier_rounds;
          if (leader) atomicAdd((int *)bar, 1);
          uint32_t spins = 0;
          while (*bar < expect && *broken == 0) {
            __nanosleep(64);
            if (++spins > (1u << 17)) *broken = 1;   // ~10-20 ms: two orders of magnitude above a round's skew
          }
        }
        const PassDesc pd = p.passes[w.pass];
        const int d = pd.d;
        const int last = 2 * d + 2;
        const int my_m0 = w.m0 + (int)rank * BLOCK_M;
        const int my_n0 = w.n0 + (int)rank * half_n;
        for (int u = pd.unit_begin; u < pd.unit_end; u++) {
          const UnitDesc ud = p.units[u];
          const int pa0 = p.plane_a[w.batch][ud.term], pb0 = p.plane_b[w.batch][ud.term];
          for (int kb = ud.kb_begin; kb < ud.kb_end; kb++) {
            const int k0 = kb * BLOCK_K;
            for (int s = 0; s <= last; s++) {
              const int t = s >> 1;
              if ((s & 1) == 0) {
                const int slot = a_it % A2_SLOTS;
                mbar_wait(emptyA(slot), ((a_it / A2_SLOTS) & 1) ^ 1);
                if (leader) mbar_expect_tx(fullA(slot), 2 * A_TILE_BYTES);  // both CTAs' halves
                tma_load_3d_2sm(smem_u32(sA + slot * A_TILE_BYTES), &map_a, leader_address(fullA(slot)), k0, my_m0, pa0 + t);
                a_it++;
              }
              if (s == 0 || (s & 1) == 1) {
                const int j = (s == 0) ? d + 1 : d - t;
                const int slot = b_it % B2_SLOTS;
                mbar_wait(emptyB(slot), ((b_it / B2_SLOTS) & 1) ^ 1);
                if (leader) mbar_expect_tx(fullB(slot), 2 * b_half_bytes);
                tma_load_3d_2sm(smem_u32(sB + slot * B2_TILE_BYTES), &map_b, leader_address(fullB(slot)), k0, my_n0, pb0 + j);
                b_it++;
              }
            }
          }
        }
      }
    }
  } else if (warp == 5) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      // D = S32, A = B = U8, K-major, N = block_n, M = 256 (cta_group::2)
      const uint32_t idesc = (2u << 4) | ((uint32_t)(p.block_n >> 3) << 17) | ((uint32_t)((2 * BLOCK_M) >> 4) << 24);
      int a_it = 0, b_it = 0;
      int a_slot = 0, b_slot = 0;
      uint32_t acc_phase = 0;
      WorkItem w;
      for (int item = 0; next_item_2cta(p, item, w); item++) {
        const PassDesc pd = p.passes[w.pass];
        const int d = pd.d;
        const int last = 2 * d + 2;
        mbar_wait(tmem_empty, acc_phase ^ 1);
        tcgen05_fence_after();
        bool first_lo = true, first_hi = true;
        for (int u = pd.unit_begin; u < pd.unit_end; u++) {
          const UnitDesc ud = p.units[u];
          for (int kb = ud.kb_begin; kb < ud.kb_end; kb++) {
            for (int s = 0; s <= last; s++) {
              const bool new_a = (s & 1) == 0;
              const bool new_b = s == 0 || (s & 1) == 1;
              const bool hi = (s & 1) == 0;
              const bool rel_a = (s & 1) == 1 || s == last;
              const bool rel_b = (s & 1) == 0;
              if (new_a) {
                a_slot = a_it % A2_SLOTS;
                mbar_wait(fullA(a_slot), (a_it / A2_SLOTS) & 1);
                a_it++;
              }
              if (new_b) {
                b_slot = b_it % B2_SLOTS;
                mbar_wait(fullB(b_slot), (b_it / B2_SLOTS) & 1);
                b_it++;
              }
              tcgen05_fence_after();
              const uint64_t adesc = make_smem_desc(smem_u32(sA + a_slot * A_TILE_BYTES));
              const uint64_t bdesc = make_smem_desc(smem_u32(sB + b_slot * B2_TILE_BYTES));
              const uint32_t tmem_d = tmem_base + (hi ? (uint32_t)MAX_BLOCK_N : 0u);
              bool &first = hi ? first_hi : first_lo;
#pragma unroll
              for (int k = 0; k < BLOCK_K / UMMA_K; k++) {
                umma_i8_2sm(tmem_d, adesc + (uint64_t)(k * UMMA_K >> 4), bdesc + (uint64_t)(k * UMMA_K >> 4), idesc,
                            (first && k == 0) ? 0u : 1u);
              }
              first = false;
              if (rel_a) umma_commit_2sm(emptyA(a_slot));
              if (rel_b) umma_commit_2sm(emptyB(b_slot));
            }
          }
        }
        umma_commit_2sm(tmem_full);
        acc_phase ^= 1;
      }
    }
  } else {
    // ===================== epilogue (both CTAs; this CTA owns rows m0 + 128 rank ..) =====================
    // TMEM hands lane r the 32 columns of row r.  Written directly, every store instruction would touch 32 rows
    // with 16 bytes each; the warp transposes through shared memory instead so that a store instruction covers one
    // row with 256 contiguous bytes (8x fewer memory transactions for the 4 GB of partial sums per launch).
    uint32_t acc_phase = 0;
    WorkItem w;
    unsigned long long *stage = (unsigned long long *)((uint8_t *)bars + 256) + (size_t)warp * 32 * STAGE_STRIDE;
    for (int item = 0; next_item_2cta(p, item, w); item++) {
      const int row0 = w.m0 + (int)rank * BLOCK_M + warp * 32;
      const int row = row0 + lane;
      const int n0 = w.n0;
      mbar_wait_relaxed(tmem_full, acc_phase);
      tcgen05_fence_after();
      unsigned long long *tile_out = p.partials + (((size_t)w.batch * p.num_passes + (size_t)w.pass) * p.M) * p.N + n0;
      unsigned long long *out = tile_out + (size_t)row * p.N;
      for (int c = 0; c < p.block_n; c += 32) {
        uint32_t lo[32], hi[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c;
        tmem_ld32(taddr, lo);
        tmem_ld32(taddr + MAX_BLOCK_N, hi);
        tmem_ld_wait();
        if (n0 + c + 32 <= p.N) {
#pragma unroll
          for (int q = 0; q < 32; q++) stage[lane * STAGE_STRIDE + q] = (unsigned long long)lo[q] + ((unsigned long long)hi[q] << 8);
          __syncwarp();
#pragma unroll 8
          for (int rr = 0; rr < 32; rr++) {  // one row per instruction: 32 lanes x 8 bytes = 256 contiguous bytes
            const unsigned long long v = stage[rr * STAGE_STRIDE + lane];
            if (row0 + rr < p.M) tile_out[(size_t)(row0 + rr) * p.N + c + lane] = v;
          }
          __syncwarp();
        } else if (row < p.M) {
#pragma unroll
          for (int q = 0; q < 32; q++) {
            if (n0 + c + q < p.N) out[c + q] = (unsigned long long)lo[q] + ((unsigned long long)hi[q] << 8);
          }
        }
      }
      tcgen05_fence_before();
      mbar_arrive_leader(tmem_empty);
      acc_phase ^= 1;
    }
  }

  tcgen05_fence_before();
  cluster_sync_all();  // neither CTA may exit (or free TMEM) while the pair still signals / reads across
  if (warp == 5) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// ---------------------------------------------------------------- limb split (operand preparation)
// planes[(term*L + i)][row][k] = byte i of A[row][k]; rows beyond the matrix are never read (TMA zero-fills
// out-of-bounds rows), columns K..Kp-1 are zero (buffer is memset before).
template <typename T>
__global__ void __launch_bounds__(256) split_a_kernel(const T *__restrict__ A, uint8_t *__restrict__ planes, int64_t M,
                                                     int64_t K, int64_t Kp) {
  constexpr int L = sizeof(T);
  // thread handles 4 consecutive k of one row
  const int64_t quads_per_row = (K + 3) / 4;
  const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= M * quads_per_row) return;
  const int64_t row = idx / quads_per_row, k = (idx % quads_per_row) * 4;
  T v[4];
#pragma unroll
  for (int q = 0; q < 4; q++) v[q] = (k + q < K) ? A[row * K + k + q] : (T)0;
#pragma unroll
  for (int i = 0; i < L; i++) {
    uint32_t w = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) w |= (uint32_t)((v[q] >> (8 * i)) & 0xff) << (8 * q);
    *(uint32_t *)(planes + ((int64_t)i * M + row) * Kp + k) = w;
  }
}

// planes[(term*L + j)][n][k] = byte j of (B0[k][n] + B1[k][n]): transposes through shared memory so both the
// global reads (n contiguous) and writes (k contiguous) move full 32-byte sectors.
template <typename T>
__global__ void __launch_bounds__(256) split_bt_kernel(const T *__restrict__ B0, const T *__restrict__ B1,
                                                      uint8_t *__restrict__ planes, int64_t K, int64_t N, int64_t Kp) {
  constexpr int L = sizeof(T);
  constexpr int PITCH = 36;  // 32 k-bytes + 4 pad: bank = 9*n mod 32 is a permutation -> conflict-free byte stores
  __shared__ __align__(16) uint8_t sm[L][32][PITCH];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // tx -> n, ty -> k (8 rows per sweep)
  const int64_t n0 = (int64_t)blockIdx.x * 32, k0 = (int64_t)blockIdx.y * 32;
#pragma unroll
  for (int it = 0; it < 4; it++) {
    const int kk = ty + 8 * it;
    const int64_t k = k0 + kk, n = n0 + tx;
    T v = 0;
    if (k < K && n < N) {
      v = B0[k * N + n];
      if (B1) v += B1[k * N + n];
    }
#pragma unroll
    for (int j = 0; j < L; j++) sm[j][tx][kk] = (uint8_t)(v >> (8 * j));
  }
  __syncthreads();
  // L * 32 * 8 words of 4 k-bytes
  for (int w = threadIdx.x; w < L * 32 * 8; w += 256) {
    const int j = w >> 8, n = (w >> 3) & 31, kq = w & 7;
    const int64_t gn = n0 + n, gk = k0 + 4 * kq;
    if (gn < N && gk < Kp) *(uint32_t *)(planes + ((int64_t)j * N + gn) * Kp + gk) = *(const uint32_t *)&sm[j][n][4 * kq];
  }
}

// C[m][n] = sum_p partial[p][m][n] << (8 d_p)   mod 2^k
struct RecombineDesc {
  int num_passes;
  int shift[64];
};
template <typename T>
__global__ void __launch_bounds__(256) recombine_kernel(const unsigned long long *__restrict__ partials, T *__restrict__ C,
                                                       int64_t MN, RecombineDesc rd) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= MN) return;
  T acc = 0;
  for (int p = 0; p < rd.num_passes; p++) acc += (T)partials[(int64_t)p * MN + i] << rd.shift[p];
  C[i] = acc;
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)ptr;
  }
  return fn;
}

int make_plane_map(mb_ctx *ctx, CUtensorMap *map, void *base, int64_t Kp, int64_t rows, int64_t planes, int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return mb_fail(ctx, MB_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t dims[3] = {(cuuint64_t)Kp, (cuuint64_t)rows, (cuuint64_t)planes};
  cuuint64_t strides[2] = {(cuuint64_t)Kp, (cuuint64_t)(Kp * rows)};
  cuuint32_t box[3] = {(cuuint32_t)BLOCK_K, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return mb_fail(ctx, MB_ERR_CUDA, "cuTensorMapEncodeTiled failed: " + std::to_string((int)r));
  return MB_OK;
}

// Pass planner: diagonal pairs (d, d+1), each cut into passes whose accumulators stay exact.
void plan_passes(int L, int kbits, int nterms, int64_t kblocks, std::vector<PassDesc> &passes, std::vector<UnitDesc> &units) {
  for (int d = 0; d < L; d += 2) {
    const bool two = d + 1 < L;
    // significant bits of C_d after the shift by 8d
    const bool may_wrap = kbits - 8 * d <= 32;
    // products accumulated per k-block into the fuller diagonal of the pair
    const int64_t per_kb = (int64_t)(two ? d + 2 : d + 1) * BLOCK_K;
    const int64_t max_kb = may_wrap ? INT64_MAX : std::max<int64_t>(1, (int64_t)ACC_LIMIT / per_kb);
    PassDesc cur = {(short)d, (short)units.size(), (short)units.size(), 0};
    int64_t used = 0;
    for (int t = 0; t < nterms; t++) {
      int64_t kb = 0;
      while (kb < kblocks) {
        if (used == max_kb) {
          cur.unit_end = (short)units.size();
          passes.push_back(cur);
          cur = {(short)d, (short)units.size(), (short)units.size(), 0};
          used = 0;
        }
        int64_t take = std::min<int64_t>(kblocks - kb, max_kb - used);
        units.push_back({t, (int)kb, (int)(kb + take)});
        kb += take;
        used += take;
      }
    }
    cur.unit_end = (short)units.size();
    passes.push_back(cur);
  }
}

bool plan_fits(mb_dtype dtype, int nterms, int64_t K, std::vector<PassDesc> &passes, std::vector<UnitDesc> &units) {
  const int L = dtype == MB_RING128 ? 16 : 8;
  const int64_t kblocks = (K + BLOCK_K - 1) / BLOCK_K;
  passes.clear();
  units.clear();
  plan_passes(L, 8 * L, nterms, kblocks, passes, units);
  return passes.size() <= MAX_PASSES && units.size() <= MAX_UNITS;
}

}  // namespace

// Co-resident 2-CTA clusters of limb_gemm_kernel_2cta on the context's device (cached per device).
static int max_active_clusters(mb_ctx *ctx, int *out) {
  static std::mutex mu;
  static int cached[64];
  std::lock_guard<std::mutex> lock(mu);
  int &c = cached[ctx->device & 63];
  if (c == 0) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(2 * (ctx->sm_count / 2)), 1, 1);
    cfg.blockDim = dim3(NUM_THREADS, 1, 1);
    cfg.dynamicSmemBytes = SMEM2_BYTES;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 2;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, (const void *)limb_gemm_kernel_2cta, &cfg);
    if (e != cudaSuccess || n <= 0) {
      (void)cudaGetLastError();
      n = ctx->sm_count / 2;   // the query is advisory; the bounded round barrier keeps an optimistic grid safe
    }
    c = std::min(n, ctx->sm_count / 2);
  }
  *out = c;
  return MB_OK;
}

// Shapes for which the tensor-pipe kernel is the better choice in automatic mode.  Measured
// (tools/gemm_crossover.py, profiles/r2_gemm_imad_vs_tensor_crossover.md): it wins on every shape it supports, from
// 128^3 (0.030 ms against 0.045 ms for the IMAD kernel on Ring64, 0.035 against 0.104 on Ring128) to 65536 x 784 x 512
// (5.6x / 9.5x), split and recombine passes included -- so the IMAD kernels serve only what the tensor-pipe kernel does
// not take (M < 128, K < 128, N < 64 or odd) and the skinny right operands.
bool k_gemm_limb_preferred(int64_t M, int64_t K, int64_t N) {
  return N >= 64 && (double)M * (double)N * (double)K >= (double)(1 << 21);
}

bool k_gemm_limb_supported(mb_dtype dtype, int64_t M, int64_t K, int64_t N) {
  if (dtype != MB_RING64 && dtype != MB_RING128) return false;
  if (M < 128 || N < 64 || K < 128) return false;
  if (N % 2) return false;                                    // 16-byte partial stores
  if (M > (1 << 30) || N > (1 << 30) || K > (1 << 22)) return false;
  std::vector<PassDesc> passes;
  std::vector<UnitDesc> units;
  return plan_fits(dtype, 2, K, passes, units);  // worst case: the two-term form used by mb_rep_dot
}

// Batched ring GEMM:  C[b] = sum_{t < nterms} A[a_idx[b][t]] * B[b_idx[b][t]]   for b < nbatch,
// where the distinct operands are limb-split once (A: nA matrices [M,K]; B: nB matrices [K,N], each optionally
// the sum of two tensors).  mb_rep_dot uses nbatch = 3, nterms = 2, nA = 3 (x_0,x_1,x_2) and nB = 6.
int k_gemm_limb_batched(mb_ctx *ctx, mb_dtype dtype, const LimbOperandA *As, int nA, const LimbOperandB *Bs, int nB,
                        const int (*a_idx)[2], const int (*b_idx)[2], int nbatch, int nterms, void *const *C, int64_t M,
                        int64_t K, int64_t N) {
  const int L = dtype == MB_RING128 ? 16 : 8;
  const int64_t Kp = (K + BLOCK_K - 1) / BLOCK_K * BLOCK_K;
  const int block_n = N >= 256 ? 256 : (N >= 128 ? 128 : 64);
  // Limb planes and partial sums live in per-context scratch: split -> GEMM -> recombine of one call must reach the
  // stream as a unit, or two host threads calling mb_dot on one context would overwrite each other's planes
  // (the reference runs every op as its own task, execution/asynchronous.rs:453-531).
  std::lock_guard<std::mutex> gemm_lock(ctx->gemm_mu);
  const bool sole_gemm_ctx = mb_gemm_ctx_register(ctx) == 1;

  std::vector<PassDesc> passes;
  std::vector<UnitDesc> units;
  if (nterms < 1 || nterms > 2) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "limb GEMM: 1 or 2 operand pairs");
  if (nbatch < 1 || nbatch > MAX_BATCH) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "limb GEMM: batch out of range");
  if ((int64_t)(nA > nB ? nA : nB) * L > 32000) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "limb GEMM: too many operands");
  if (!plan_fits(dtype, nterms, K, passes, units))
    return mb_fail(ctx, MB_ERR_KERNEL, "limb GEMM: K too large for the pass table");
  // longest passes first (static LPT schedule); cost = MMA steps
  auto cost = [&](const PassDesc &pd) {
    int64_t kb = 0;
    for (int u = pd.unit_begin; u < pd.unit_end; u++) kb += units[u].kb_end - units[u].kb_begin;
    return kb * (pd.d + 1 < L ? 2 * pd.d + 3 : pd.d + 1);
  };
  std::stable_sort(passes.begin(), passes.end(), [&](const PassDesc &x, const PassDesc &y) { return cost(x) > cost(y); });

  const size_t a_plane_bytes = (size_t)L * M * Kp, b_plane_bytes = (size_t)L * N * Kp;
  const size_t a_bytes = (size_t)nA * a_plane_bytes;
  const size_t b_bytes = (size_t)nB * b_plane_bytes;
  const size_t part_bytes = (size_t)nbatch * passes.size() * (size_t)M * N * 8;
  void *dA, *dB, *dP;
  MB_CHECK(mb_scratch(ctx, 0, a_bytes, &dA));
  MB_CHECK(mb_scratch(ctx, 1, b_bytes, &dB));
  MB_CHECK(mb_scratch(ctx, 2, part_bytes, &dP));

  // operand preparation: every distinct operand is split into byte planes once
  if (Kp != K) MB_CUDA(ctx, cudaMemsetAsync(dA, 0, a_bytes, ctx->stream));
  const int64_t quads = M * ((K + 3) / 4);
  dim3 gb((unsigned)((N + 31) / 32), (unsigned)((Kp + 31) / 32));
  for (int i = 0; i < nA; i++) {
    uint8_t *pa = (uint8_t *)dA + (size_t)i * a_plane_bytes;
    if (dtype == MB_RING128)
      split_a_kernel<u128><<<(unsigned)((quads + 255) / 256), 256, 0, ctx->stream>>>((const u128 *)As[i].ptr, pa, M, K, Kp);
    else
      split_a_kernel<uint64_t><<<(unsigned)((quads + 255) / 256), 256, 0, ctx->stream>>>((const uint64_t *)As[i].ptr, pa, M, K, Kp);
    MB_LAUNCH_CHECK_W(ctx, "limb_split", 2.0 * L * M * K, 0, 0);
  }
  for (int i = 0; i < nB; i++) {
    uint8_t *pb = (uint8_t *)dB + (size_t)i * b_plane_bytes;
    if (dtype == MB_RING128)
      split_bt_kernel<u128><<<gb, 256, 0, ctx->stream>>>((const u128 *)Bs[i].b0, (const u128 *)Bs[i].b1, pb, K, N, Kp);
    else
      split_bt_kernel<uint64_t><<<gb, 256, 0, ctx->stream>>>((const uint64_t *)Bs[i].b0, (const uint64_t *)Bs[i].b1, pb, K, N, Kp);
    MB_LAUNCH_CHECK_W(ctx, "limb_split", (Bs[i].b1 ? 3.0 : 2.0) * L * N * K, 0, 0);
  }

  const bool use_pair = ctx->gemm_pair_mode != 1 && M >= 2 * BLOCK_M && N >= 128;
  CUtensorMap map_a, map_b;
  MB_CHECK(make_plane_map(ctx, &map_a, dA, Kp, M, (int64_t)nA * L, BLOCK_M));
  MB_CHECK(make_plane_map(ctx, &map_b, dB, Kp, N, (int64_t)nB * L, use_pair ? block_n / 2 : block_n));

  GemmParams gp;
  memset(&gp, 0, sizeof(gp));
  memcpy(gp.passes, passes.data(), passes.size() * sizeof(PassDesc));
  memcpy(gp.units, units.data(), units.size() * sizeof(UnitDesc));
  for (int b = 0; b < nbatch; b++)
    for (int t = 0; t < nterms; t++) {
      if (a_idx[b][t] < 0 || a_idx[b][t] >= nA || b_idx[b][t] < 0 || b_idx[b][t] >= nB)
        return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "limb GEMM: operand index out of range");
      gp.plane_a[b][t] = (short)(a_idx[b][t] * L);
      gp.plane_b[b][t] = (short)(b_idx[b][t] * L);
    }
  gp.num_passes = (int)passes.size();
  gp.nbatch = nbatch;
  gp.L = L;
  gp.M = (int)M;
  gp.N = (int)N;
  gp.block_n = block_n;
  // CTA pairs (cta_group::2, 256-row tiles) whenever the problem is large enough; L = 8 / 16 is even, so
  // every pass is a full diagonal pair as the 2-CTA kernel assumes
  const bool pair = use_pair;
  const int tile_m = pair ? 2 * BLOCK_M : BLOCK_M;
  gp.tiles_m = (int)((M + tile_m - 1) / tile_m);
  gp.tiles_n = (int)((N + block_n - 1) / block_n);
  const int64_t total = (int64_t)gp.num_passes * nbatch * gp.tiles_m * gp.tiles_n;
  if (total > INT32_MAX) return mb_fail(ctx, MB_ERR_KERNEL, "limb GEMM: too many work items");
  gp.total_items = (int)total;
  gp.partials = (unsigned long long *)dP;

  // persistent grid: as many clusters as can be resident at once (per-GPC cluster placement, ~200 KB of shared memory
  // and all 512 TMEM columns per CTA make that fewer than sm_count / 2 on some parts)
  int clusters = 0;
  if (pair) {
    MB_CUDA(ctx, cudaFuncSetAttribute(limb_gemm_kernel_2cta, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM2_BYTES));
    MB_CHECK(max_active_clusters(ctx, &clusters));
    clusters = (int)std::min<int64_t>(total, clusters);
  }
  gp.round_barrier = nullptr;
  gp.barrier_rounds = 0;
  // the round barrier only pays (and is only safe to lean on) when this context has the GPU's tensor pipes to itself:
  // with several contexts running limb GEMMs on one device (the in-process three-party mode) clusters of different
  // launches hold each other's SMs, so those launches run free
  if (pair && ctx->gemm_round_barrier && sole_gemm_ctx) {
    const int64_t rounds = (total + clusters - 1) / clusters;
    void *dR;
    MB_CHECK(mb_scratch(ctx, 3, (size_t)(rounds + 1) * sizeof(int), &dR));
    MB_CUDA(ctx, cudaMemsetAsync(dR, 0, (size_t)(rounds + 1) * sizeof(int), ctx->stream));
    gp.round_barrier = (int *)dR;
    gp.barrier_rounds = (int)rounds;
  }
  cudaEvent_t prof_start;
  MB_CHECK(mb_prof_begin(ctx, &prof_start));
  if (pair) {
    limb_gemm_kernel_2cta<<<2 * clusters, NUM_THREADS, SMEM2_BYTES, ctx->stream>>>(map_a, map_b, gp);
  } else {
    MB_CUDA(ctx, cudaFuncSetAttribute(limb_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    const int grid = (int)std::min<int64_t>(total, ctx->sm_count);
    limb_gemm_kernel<<<grid, NUM_THREADS, SMEM_BYTES, ctx->stream>>>(map_a, map_b, gp);
  }
  MB_LAUNCH_CHECK_W(ctx, "limb_gemm", (double)a_bytes + (double)b_bytes + (double)part_bytes, 0,
                    (double)nbatch * nterms * 2.0 * M * N * Kp * (L * (L + 1) / 2));
  MB_CHECK(mb_prof_end(ctx, prof_start));

  RecombineDesc rd;
  rd.num_passes = (int)passes.size();
  for (size_t i = 0; i < passes.size(); i++) rd.shift[i] = 8 * passes[i].d;
  const int64_t MN = M * N;
  for (int b = 0; b < nbatch; b++) {
    const unsigned long long *pp = (const unsigned long long *)dP + (size_t)b * passes.size() * MN;
    if (dtype == MB_RING128)
      recombine_kernel<u128><<<(unsigned)((MN + 255) / 256), 256, 0, ctx->stream>>>(pp, (u128 *)C[b], MN, rd);
    else
      recombine_kernel<uint64_t><<<(unsigned)((MN + 255) / 256), 256, 0, ctx->stream>>>(pp, (uint64_t *)C[b], MN, rd);
    MB_LAUNCH_CHECK_W(ctx, "limb_recombine", (double)MN * (8.0 * passes.size() + L), 0, 0);
  }
  return MB_OK;
}

int k_gemm_limb(mb_ctx *ctx, mb_dtype dtype, const GemmTerm *terms, int nterms, void *C, int64_t M, int64_t K, int64_t N) {
  if (nterms < 1 || nterms > 2) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "limb GEMM: 1 or 2 operand pairs");
  LimbOperandA As[2];
  LimbOperandB Bs[2];
  int a_idx[1][2] = {{0, 1}}, b_idx[1][2] = {{0, 1}};
  for (int t = 0; t < nterms; t++) {
    As[t].ptr = terms[t].A;
    Bs[t].b0 = terms[t].B0;
    Bs[t].b1 = terms[t].B1;
  }
  void *Cs[1] = {C};
  return k_gemm_limb_batched(ctx, dtype, As, nterms, Bs, nterms, a_idx, b_idx, 1, nterms, Cs, M, K, N);
}

// =====================================================================================
// mb_dot_noalloc: the tensor-pipe ring GEMM for callers that must not allocate.
//
// Everything k_gemm_limb keeps in per-context scratch buffers (the byte-limb planes of both
// operands and the per-pass partial sums) lives here in ONE fixed array that is part of the
// library image.  The array exists once per GPU that loads the library and stays resident
// until the process exits, which is why its size is kept modest.
// =====================================================================================
namespace {

constexpr size_t NOALLOC_SCRATCH_BYTES = 8u << 20;  // 8 MiB
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
unsigned char *g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of g_noalloc_scratch, per device (0 = not set up)

// In host code the identifier g_noalloc_scratch is only a handle for the runtime: its device address has to
// be asked for, once per device and with that device current.
int noalloc_scratch_address(mb_ctx *ctx, unsigned char **out) {
  if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_dot_noalloc: device index out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[ctx->device]) {
    MB_CUDA(ctx, cudaSetDevice(ctx->device));
    void *p = nullptr;
    MB_CUDA(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) return mb_fail(ctx, MB_ERR_CUDA, "mb_dot_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[ctx->device] = (unsigned char *)p;
  }
  *out = g_noalloc_ptr[ctx->device];
  return MB_OK;
}

// plan_passes without heap containers: fills the kernel parameter block directly.
bool plan_passes_fixed(int L, int kbits, int64_t kblocks, GemmParams &gp) {
  int np = 0, nu = 0;
  for (int d = 0; d < L; d += 2) {
    const bool two = d + 1 < L;
    const bool may_wrap = kbits - 8 * d <= 32;
    const int64_t per_kb = (int64_t)(two ? d + 2 : d + 1) * BLOCK_K;
    const int64_t max_kb = may_wrap ? INT64_MAX : std::max<int64_t>(1, (int64_t)ACC_LIMIT / per_kb);
    int64_t kb = 0;
    while (kb < kblocks) {
      if (np >= MAX_PASSES || nu >= MAX_UNITS) return false;
      const int64_t take = std::min<int64_t>(kblocks - kb, max_kb);
      gp.units[nu] = {0, (int)kb, (int)(kb + take)};
      gp.passes[np] = {(short)d, (short)nu, (short)(nu + 1), 0};
      nu++;
      np++;
      kb += take;
    }
  }
  gp.num_passes = np;
  return true;
}

inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" int mb_dot_noalloc(mb_ctx *ctx, mb_dtype dtype, const void *a_dev, const void *b_dev, void *c_dev, int64_t M,
                              int64_t K, int64_t N) {
  if (!ctx) return mb_fail(nullptr, MB_ERR_INVALID_ARGUMENT, "mb_dot_noalloc: NULL context");
  if (!a_dev || !b_dev || !c_dev) return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_dot_noalloc: NULL operand pointer");
  if (dtype != MB_RING64 && dtype != MB_RING128)
    return mb_fail(ctx, MB_ERR_TYPE_MISMATCH, "mb_dot_noalloc: ring64 or ring128 operands only");
  if (M < 128 || N < 64 || K < 128 || (N % 2) != 0)
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_dot_noalloc: needs M >= 128, K >= 128, N >= 64 and N even");
  MB_CUDA(ctx, cudaSetDevice(ctx->device));

  const int L = dtype == MB_RING128 ? 16 : 8;
  const int64_t Kp = (K + BLOCK_K - 1) / BLOCK_K * BLOCK_K;
  const int block_n = N >= 256 ? 256 : (N >= 128 ? 128 : 64);

  GemmParams gp;
  memset(&gp, 0, sizeof(gp));
  if (!plan_passes_fixed(L, 8 * L, Kp / BLOCK_K, gp))
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT, "mb_dot_noalloc: K too large for the pass table");

  // Carve the static array: [A limb planes | B limb planes | partial sums], each 1024-byte aligned.
  const size_t a_bytes = round_up((size_t)L * M * Kp, 1024);
  const size_t b_bytes = round_up((size_t)L * N * Kp, 1024);
  const size_t part_bytes = (size_t)gp.num_passes * M * N * 8;
  const size_t need = a_bytes + b_bytes + part_bytes;
  if (need > NOALLOC_SCRATCH_BYTES)
    return mb_fail(ctx, MB_ERR_INVALID_ARGUMENT,
                   "mb_dot_noalloc: this shape needs " + std::to_string(need) +
                       " bytes of scratch but the built-in scratch array holds " +
                       std::to_string(NOALLOC_SCRATCH_BYTES) + " bytes; use mb_dot for larger problems");

  unsigned char *scratch = nullptr;
  MB_CHECK(noalloc_scratch_address(ctx, &scratch));
  uint8_t *dA = scratch, *dB = scratch + a_bytes;
  unsigned long long *dP = (unsigned long long *)(scratch + a_bytes + b_bytes);

  // The array still holds whatever the previous call left behind.  The only bytes that are read without
  // being written first are the K..Kp padding columns of the A planes, so those planes are zeroed when K is
  // not a multiple of the 128-byte tile (split_bt_kernel writes its own padding; partials are fully overwritten).
  if (Kp != K) MB_CUDA(ctx, cudaMemsetAsync(dA, 0, (size_t)L * M * Kp, ctx->stream));

  const int64_t quads = M * ((K + 3) / 4);
  dim3 gb((unsigned)((N + 31) / 32), (unsigned)((Kp + 31) / 32));
  if (dtype == MB_RING128) {
    split_a_kernel<u128><<<(unsigned)((quads + 255) / 256), 256, 0, ctx->stream>>>((const u128 *)a_dev, dA, M, K, Kp);
    MB_LAUNCH_CHECK(ctx);
    split_bt_kernel<u128><<<gb, 256, 0, ctx->stream>>>((const u128 *)b_dev, (const u128 *)nullptr, dB, K, N, Kp);
    MB_LAUNCH_CHECK(ctx);
  } else {
    split_a_kernel<uint64_t><<<(unsigned)((quads + 255) / 256), 256, 0, ctx->stream>>>((const uint64_t *)a_dev, dA, M, K, Kp);
    MB_LAUNCH_CHECK(ctx);
    split_bt_kernel<uint64_t><<<gb, 256, 0, ctx->stream>>>((const uint64_t *)b_dev, (const uint64_t *)nullptr, dB, K, N, Kp);
    MB_LAUNCH_CHECK(ctx);
  }

  CUtensorMap map_a, map_b;
  MB_CHECK(make_plane_map(ctx, &map_a, dA, Kp, M, L, BLOCK_M));
  MB_CHECK(make_plane_map(ctx, &map_b, dB, Kp, N, L, block_n));

  gp.nbatch = 1;  // plane_a / plane_b stay 0: one operand pair
  gp.L = L;
  gp.M = (int)M;
  gp.N = (int)N;
  gp.block_n = block_n;
  gp.tiles_m = (int)((M + BLOCK_M - 1) / BLOCK_M);
  gp.tiles_n = (int)((N + block_n - 1) / block_n);
  gp.total_items = gp.num_passes * gp.tiles_m * gp.tiles_n;
  gp.partials = dP;

  MB_CUDA(ctx, cudaFuncSetAttribute(limb_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  const int grid = std::min(gp.total_items, ctx->sm_count);
  limb_gemm_kernel<<<grid, NUM_THREADS, SMEM_BYTES, ctx->stream>>>(map_a, map_b, gp);
  MB_LAUNCH_CHECK(ctx);

  RecombineDesc rd;
  rd.num_passes = gp.num_passes;
  for (int i = 0; i < gp.num_passes; i++) rd.shift[i] = 8 * gp.passes[i].d;
  const int64_t MN = M * N;
  if (dtype == MB_RING128)
    recombine_kernel<u128><<<(unsigned)((MN + 255) / 256), 256, 0, ctx->stream>>>(dP, (u128 *)c_dev, MN, rd);
  else
    recombine_kernel<uint64_t><<<(unsigned)((MN + 255) / 256), 256, 0, ctx->stream>>>(dP, (uint64_t *)c_dev, MN, rd);
  MB_LAUNCH_CHECK(ctx);
  return MB_OK;
}
