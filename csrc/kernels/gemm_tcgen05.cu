// K6: tcgen05 / TMEM GEMMs fed by TMA, with fused epilogues.
//
// Replaces the cuBLAS calls behind nn.Linear forward / backward in the reference's training
// loop (reference python/raydp/torch/estimator.py:276-286 -> model(inputs), loss.backward()).
//
//  * rdp_gemm_tn   : C[M,N] = A[M,K] . B[N,K]^T  (both operands K-major: forward with W[N,K],
//                    dgrad with a pre-transposed weight).  Persistent, warp-specialised:
//                    warp 0 = TMA producer, warp 1 = MMA issuer (one elected thread issues
//                    tcgen05.mma), warps 2-5 = epilogue (tcgen05.ld from a double-buffered TMEM
//                    accumulator so the epilogue of tile i overlaps the MMAs of tile i+1).
//                    Epilogue options: +bias, ReLU, ReLU-mask by an aux matrix, bf16/fp32 store,
//                    and per-column batch statistics (sum & sum-of-squares, or sum & sum(v*aux))
//                    accumulated in shared memory and flushed once per CTA -- this is what lets
//                    BatchNorm's batch statistics ride on the GEMM instead of re-reading the
//                    activations.
//  * rdp_gemm_wgrad: D[Mw,Nw] += sum_r G[r,Mw] . X[r,Nw]  (reduction over the batch dimension,
//                    both operands MN-major straight out of the row-major activations, split
//                    over CTAs along the batch, fp32 atomics into D).
#include "common.cuh"
#include <stdlib.h>
#include "raydp_b200.h"

#include <mutex>

namespace rdp {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;          // 64 bf16 = 128 B = one swizzle row
constexpr int kMaxNTile = 256;
constexpr int kStages = 4;
constexpr int kEpiGroups = 3;        // column groups: 3 x 4 quadrant warps = 12 epilogue warps
constexpr int kGemmThreads = 64 + kEpiGroups * 128;   // TMA warp + MMA warp + epilogue warps
constexpr int kStatsCols = 1024;

struct GemmTnParams {
  int M, N, K;
  int n_tile;
  int num_m_tiles, num_n_tiles, num_k_blocks;
  const float* bias;
  int relu;
  void* out;
  int out_f32;
  int ldc;
  const __nv_bfloat16* aux;
  int ld_aux;
  int aux_relu_mask;
  int stats_mode;   // 0 none | 1: sum(v), sum(v*v) | 2: sum(v), sum(v*aux)
  float* stats;     // [2, N]
  // fast kernel only: second statistics row (default stats + N; null = not wanted) and the fused
  // BatchNorm+ReLU backward of the layer below (see rdp_gemm_tn_bnbwd): with aux = that layer's ReLU
  // output r, the stored value is (r > 0) ? alpha[n] * acc + beta[n] * r + delta[n] : 0
  float* stats1;
  const float* bn_alpha;
  const float* bn_beta;
  const float* bn_delta;
  long long* dbg;   // diagnostics (RDP_GEMM_DBG=1): per-tile timestamps of CTA 0, else null
};

#define RDP_DBG_STAMP(slot)                                                                         \
  do {                                                                                              \
    if (p.dbg != nullptr && blockIdx.x == 0 && dbg_tile < 32) {                                     \
      long long _t;                                                                                 \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t));                                        \
      p.dbg[dbg_tile * 16 + (slot)] = _t;                                                           \
    }                                                                                               \
  } while (0)

constexpr int kBnCoefCols = 256;     // widest layer the fused BN-backward epilogue keeps coefficients for

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tn_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const GemmTnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve shared memory: [stage tiles][barriers][tmem ptr][stats]
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t a_bytes = kBlockM * kBlockK * 2;                 // 16 KB
  const uint32_t b_bytes = uint32_t(p.n_tile) * kBlockK * 2;      // n_tile * 128 B
  const uint32_t stage_bytes = a_bytes + ((b_bytes + 1023) & ~1023u);
  uint8_t* tiles = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + kStages * stage_bytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* s_stats = reinterpret_cast<float*>(tmem_ptr + 4);        // [2][kStatsCols]
  float* s_bias = s_stats + 2 * kStatsCols;                       // [kStatsCols]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (p.stats_mode) {
    for (int i = threadIdx.x; i < 2 * kStatsCols; i += blockDim.x) s_stats[i] = 0.f;
  }
  const bool bias_in_smem = p.bias != nullptr && p.N <= kStatsCols;
  if (bias_in_smem) {
    for (int i = threadIdx.x; i < p.N; i += blockDim.x) s_bias[i] = p.bias[i];
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], kEpiGroups * 128);
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int m_blk = t / p.num_n_tiles, n_blk = t % p.num_n_tiles;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = tiles + stage * stage_bytes;
          uint8_t* b_dst = a_dst + a_bytes;
          mbar_expect_tx(&full_bar[stage], a_bytes + b_bytes);
          tma_load_2d(a_dst, &tmap_a, &full_bar[stage], kb * kBlockK, m_blk * kBlockM);
          tma_load_2d(b_dst, &tmap_b, &full_bar[stage], kb * kBlockK, n_blk * p.n_tile);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kBlockM, uint32_t(p.n_tile), false, false);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + uint32_t(acc) * kMaxNTile;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(tiles + stage * stage_bytes);
          const uint32_t b_addr = a_addr + a_bytes;
          const uint64_t a_desc = umma_smem_desc(a_addr, 16, 1024);
          const uint64_t b_desc = umma_smem_desc(b_addr, 16, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // advance 16 elements (32 B) along K inside the 128-byte swizzle row
            umma_bf16(d_tmem, a_desc + uint64_t(k * 2), b_desc + uint64_t(k * 2), idesc,
                      (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);           // frees the smem slot when the MMAs retire
          if (kb == p.num_k_blocks - 1) umma_commit(&tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue warps (2..13) =====================
    // 12 warps: the SM's four schedulers each get three epilogue warps, which is what hides the
    // TMEM-load / shuffle / store latencies (one warp per scheduler was issue-latency bound).
    const int quad = warp & 3;                       // TMEM lane quadrant this warp may read
    const int group = (warp - 2) >> 2;               // which 32-column chunks this warp owns
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int m_blk = t / p.num_n_tiles, n_blk = t % p.num_n_tiles;
      const int row = m_blk * kBlockM + quad * 32 + lane;
      const bool row_ok = row < p.M;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (uint32_t(quad * 32) << 16) + uint32_t(acc) * kMaxNTile;
      const int n0 = n_blk * p.n_tile;
      for (int c = group * 32; c < p.n_tile; c += kEpiGroups * 32) {
        const int col0 = n0 + c;
        if (col0 >= p.N) break;
        const int ncols = min(32, min(p.n_tile - c, p.N - col0));   // 16 or 32
        uint32_t raw[32];
        tmem_ld_32x32(t_row + uint32_t(c), raw);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
        if (bias_in_smem) {
          const float4* b4 = reinterpret_cast<const float4*>(s_bias + col0);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            if (q * 4 < ncols) {
              const float4 bb = b4[q];
              v[q * 4] += bb.x; v[q * 4 + 1] += bb.y; v[q * 4 + 2] += bb.z; v[q * 4 + 3] += bb.w;
            }
          }
        } else if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += (j < ncols) ? __ldg(p.bias + col0 + j) : 0.f;
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        float aux[32];
        const bool need_aux = (p.aux != nullptr) && (p.aux_relu_mask || p.stats_mode == 2);
        if (need_aux) {
          if (row_ok) {
            const uint4* ap = reinterpret_cast<const uint4*>(p.aux + size_t(row) * p.ld_aux + col0);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (q * 8 < ncols) {
                uint4 u = __ldg(ap + q);
                float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y);
                float2 f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
                aux[q * 8 + 0] = f0.x; aux[q * 8 + 1] = f0.y; aux[q * 8 + 2] = f1.x; aux[q * 8 + 3] = f1.y;
                aux[q * 8 + 4] = f2.x; aux[q * 8 + 5] = f2.y; aux[q * 8 + 6] = f3.x; aux[q * 8 + 7] = f3.y;
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) aux[q * 8 + j] = 0.f;
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) aux[j] = 0.f;
          }
          if (p.aux_relu_mask) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = aux[j] > 0.f ? v[j] : 0.f;
          }
        }
        if (!row_ok) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
        // ---- store ----
        uint32_t packed[16];
        if (!p.out_f32) {
#pragma unroll
          for (int i = 0; i < 16; ++i) packed[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
        }
        if (row_ok) {
          if (p.out_f32) {
            float4* op = reinterpret_cast<float4*>(static_cast<float*>(p.out) + size_t(row) * p.ldc + col0);
#pragma unroll
            for (int q = 0; q < 8; ++q)
              if (q * 4 < ncols) op[q] = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
          } else {
            uint4* op = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + size_t(row) * p.ldc + col0);
#pragma unroll
            for (int q = 0; q < 4; ++q)
              if (q * 8 < ncols)
                op[q] = make_uint4(packed[q * 4], packed[q * 4 + 1], packed[q * 4 + 2], packed[q * 4 + 3]);
          }
        }
        // ---- per-column batch statistics ----
        if (p.stats_mode) {
          float w[32];
          if (p.stats_mode == 1) {
            // statistics of what the next layer will read: the bf16-rounded activation
            if (!p.out_f32) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {      // bf16 -> fp32 is a shift / mask of the packed pair
                v[2 * i] = __uint_as_float(packed[i] << 16);
                v[2 * i + 1] = __uint_as_float(packed[i] & 0xffff0000u);
              }
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) w[j] = v[j] * v[j];
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) w[j] = v[j] * aux[j];
          }
          const float s0 = warp_transpose_reduce32(v, lane);
          const float s1 = warp_transpose_reduce32(w, lane);
          if (lane < ncols) {
            const int col = col0 + lane;
            if (col < kStatsCols) {
              atomicAdd(&s_stats[col], s0);
              atomicAdd(&s_stats[kStatsCols + col], s1);
            } else {
              atomicAdd(&p.stats[col], s0);
              atomicAdd(&p.stats[p.N + col], s1);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (p.stats_mode) {
    const int ncol = min(p.N, kStatsCols);
    for (int i = threadIdx.x; i < ncol; i += blockDim.x) {
      const float a = s_stats[i], b = s_stats[kStatsCols + i];
      if (a != 0.f) atomicAdd(&p.stats[i], a);
      if (b != 0.f) atomicAdd(&p.stats[p.N + i], b);
    }
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// ----------------------------------------------------------------------------------------------
// Specialised fast path (bf16 output, N % 64 == 0, N <= kStatsCols).
//
// The generic kernel's epilogue is row-per-thread: every 16-byte store / aux load of a warp touches
// 32 different 128-byte lines, and the L1 wavefronts of those accesses -- not DRAM -- bound it
// (ncu: dram 2 %, long/short scoreboard stalls).  Here every byte moves by TMA:
//   * the aux tile (BatchNorm input z for the backward statistics / ReLU mask) is loaded by the
//     producer warp into a swizzled smem tile,
//   * the epilogue warps write the bf16 output tile into a double-buffered swizzled smem tile and
//     one thread issues `cp.async.bulk.tensor` stores (fully coalesced, asynchronous),
//   * 16 epilogue warps (4 per SM scheduler), epilogue compiled per (statistics mode, aux use),
//   * the per-column batch statistics are computed BY THE TENSOR CORES: the staged bf16 output
//     tile is exactly an MN-major UMMA operand, so the MMA warp issues two extra accumulating
//     MMAs per tile, D_s += Ones . C (every row = column sums) and D_g += C^T . C (mode 1) or
//     C^T . X (mode 2), whose diagonals are sum(v), sum(v^2) / sum(v * aux).  They accumulate in
//     TMEM across all tiles of the CTA and are read out once at the end, which removes the
//     shuffle-transpose reduction (2/3 of the epilogue's instructions) altogether.
// ----------------------------------------------------------------------------------------------
constexpr int kFastGroups = 4;
constexpr int kFastThreads = 64 + kFastGroups * 128;
constexpr int kFastNTile = 128;
constexpr uint32_t kChunkTileBytes = 128 * 128;      // one [128 rows x 64 cols] bf16 swizzled block

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kPending) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void epi_barrier() {
  asm volatile("bar.sync 1, %0;" ::"n"(kFastGroups * 128) : "memory");
}

template <int kStats, bool kAux>
__global__ void __launch_bounds__(kFastThreads, 1)
gemm_tn_fast_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_x,
                    const GemmTnParams p) {
  constexpr int kS = kAux ? 2 : 3;                    // operand pipeline depth
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t a_bytes = kBlockM * kBlockK * 2;                      // 16 KB
  const uint32_t b_bytes = uint32_t(p.n_tile) * kBlockK * 2;           // <= 16 KB
  const uint32_t stage_bytes = a_bytes + kFastNTile * kBlockK * 2;     // 32 KB
  const int n_chunks64 = p.n_tile / 64;                                // 1 or 2
  const uint32_t c_bytes = kChunkTileBytes * 2;                        // one staging buffer (<= 128 cols)
  uint8_t* tiles = smem;
  uint8_t* c_stage = tiles + kS * stage_bytes;                         // 2 buffers
  uint8_t* x_stage = c_stage + 2 * c_bytes;                            // 2 buffers (kAux only)
  uint8_t* ones = x_stage + (kAux ? 2 * c_bytes : 0);                  // [128 x 128 B] of bf16 1.0 (kStats)
  uint8_t* tail = ones + (kStats ? kChunkTileBytes : 0);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + kS;
  uint64_t* tmem_full = empty_bar + kS;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* x_full = tmem_empty + 2;
  uint64_t* x_empty = x_full + 2;
  uint64_t* c_ready = x_empty + 2;       // epilogue -> MMA warp: staged tile complete
  uint64_t* stats_done = c_ready + 2;    // MMA warp -> epilogue: statistics MMAs have read the tile
  uint64_t* stats_final = stats_done + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(stats_final + 1);
  float* s_bias = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(tmem_ptr + 4) + 15) & ~uintptr_t(15));
  float* s_coef = s_bias + kStatsCols;                                 // [3][kBnCoefCols] (kAux only)
  const bool bn_bwd = kAux && p.bn_alpha != nullptr;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_trigger();                     // the next kernel of the stream may start its own prologue
  if (kStats) {
    uint32_t* o32 = reinterpret_cast<uint32_t*>(ones);
    for (int i = threadIdx.x; i < int(kChunkTileBytes / 4); i += blockDim.x) o32[i] = 0x3F803F80u;
    fence_async_smem();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_c);
    if (kAux) tma_prefetch_desc(&tmap_x);
    for (int s = 0; s < kS; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], kFastGroups * 128);
      mbar_init(&x_full[a], 1);
      mbar_init(&x_empty[a], kStats == 2 ? 1 : kFastGroups * 128);
      mbar_init(&c_ready[a], 1);
      mbar_init(&stats_done[a], 1);
    }
    mbar_init(stats_final, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<kStats ? 512 : 256>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // Everything above touched only shared memory, TMEM and kernel parameters: it overlaps the tail of the previous
  // kernel.  From here on global memory is read (operands, bias, BN constants produced by that kernel).
  pdl_wait();
  if (bn_bwd) {
    for (int i = threadIdx.x; i < p.N; i += blockDim.x) {
      s_coef[i] = p.bn_alpha[i];
      s_coef[kBnCoefCols + i] = p.bn_beta[i];
      s_coef[2 * kBnCoefCols + i] = p.bn_delta[i];
    }
  }
  for (int i = threadIdx.x; i < p.N; i += blockDim.x) s_bias[i] = p.bias ? p.bias[i] : 0.f;
  __syncthreads();
  const uint32_t tmem_base = *tmem_ptr;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;

  if (warp == 0) {
    // ===================== TMA producer: operands + aux tile =====================
    if (lane == 0) {
      int stage = 0, xb = 0;
      uint32_t phase = 0, x_phase = 0;
      int dbg_tile = -1;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        ++dbg_tile;
        const int m_blk = t / p.num_n_tiles, n_blk = t % p.num_n_tiles;
        RDP_DBG_STAMP(0);                     // producer reaches the tile
        // operands first: the MMA of this tile must never wait behind the aux buffer, whose release
        // depends on the epilogue + statistics MMAs of two tiles ago
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = tiles + stage * stage_bytes;
          mbar_expect_tx(&full_bar[stage], a_bytes + b_bytes);
          tma_load_2d(a_dst, &tmap_a, &full_bar[stage], kb * kBlockK, m_blk * kBlockM);
          tma_load_2d(a_dst + a_bytes, &tmap_b, &full_bar[stage], kb * kBlockK, n_blk * p.n_tile);
          if (++stage == kS) { stage = 0; phase ^= 1; }
        }
        RDP_DBG_STAMP(1);                     // operand loads issued
        if (kAux) {
          mbar_wait(&x_empty[xb], x_phase ^ 1);
          mbar_expect_tx(&x_full[xb], uint32_t(n_chunks64) * kChunkTileBytes);
          for (int ch = 0; ch < n_chunks64; ++ch)
            tma_load_2d(x_stage + xb * c_bytes + ch * kChunkTileBytes, &tmap_x, &x_full[xb],
                        n_blk * p.n_tile + ch * 64, m_blk * kBlockM);
          if (++xb == 2) { xb = 0; x_phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(kBlockM, uint32_t(p.n_tile), false, false);
      // statistics MMAs: D_s = Ones(K-major) . C(MN-major) ; D_g = C^T(MN-major) . {C | X}(MN-major)
      const uint32_t idesc_s = umma_idesc_bf16(128, uint32_t(p.n_tile), false, true);
      const uint32_t idesc_g = umma_idesc_bf16(128, uint32_t(p.n_tile), true, true);
      const uint32_t d_sum = tmem_base + 2 * kFastNTile, d_gram = tmem_base + 3 * kFastNTile;
      const uint64_t ones_desc = umma_smem_desc(smem_u32(ones), 16, 1024);
      auto stats_mma = [&](int sacc, uint32_t sphase, bool first) {
        mbar_wait(&c_ready[sacc], sphase);
        tc_fence_after();
        const uint32_t c_addr = smem_u32(c_stage + sacc * c_bytes);
        const uint32_t o_addr = (kStats == 2) ? smem_u32(x_stage + sacc * c_bytes) : c_addr;
#pragma unroll
        for (int k = 0; k < 128 / 16; ++k) {
          // 16 tile rows per MMA = 2048 B down the swizzled block; 64-column chunks are 16 KB apart
          const uint64_t c_desc = umma_smem_desc(c_addr + k * 2048, kChunkTileBytes, 1024);
          const uint64_t o_desc = umma_smem_desc(o_addr + k * 2048, kChunkTileBytes, 1024);
          umma_bf16(d_sum, ones_desc, c_desc, idesc_s, !(first && k == 0));
          if (kStats != 3) umma_bf16(d_gram, c_desc, o_desc, idesc_g, !(first && k == 0));
        }
        umma_commit(&stats_done[sacc]);
        if (kStats == 2) umma_commit(&x_empty[sacc]);
      };
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      int done = 0;
      int dbg_tile = -1;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        ++dbg_tile;
        RDP_DBG_STAMP(2);                     // MMA warp reaches the tile
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        RDP_DBG_STAMP(3);                     // accumulator free
        const uint32_t d_tmem = tmem_base + uint32_t(acc) * kFastNTile;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(tiles + stage * stage_bytes);
          const uint64_t a_desc = umma_smem_desc(a_addr, 16, 1024);
          const uint64_t b_desc = umma_smem_desc(a_addr + a_bytes, 16, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)
            umma_bf16(d_tmem, a_desc + uint64_t(k * 2), b_desc + uint64_t(k * 2), idesc, (kb | k) != 0);
          umma_commit(&empty_bar[stage]);
          if (kb == p.num_k_blocks - 1) umma_commit(&tmem_full[acc]);
          if (++stage == kS) { stage = 0; phase ^= 1; }
        }
        RDP_DBG_STAMP(4);                     // operands arrived, MMAs issued
        if (kStats && done > 0) {
          // statistics of the previous tile, whose staged output the epilogue has finished by now
          const int pacc = acc ^ 1;
          const uint32_t pphase = (pacc == 1) ? acc_phase ^ 1 : acc_phase;   // phase of tile done-1
          stats_mma(pacc, (acc == 0) ? (acc_phase ^ 1) : acc_phase, done == 1);
          (void)pphase;
        }
        ++done;
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (kStats && done > 0) {
        const int pacc = acc ^ 1;
        stats_mma(pacc, (acc == 0) ? (acc_phase ^ 1) : acc_phase, done == 1);
        umma_commit(stats_final);
      }
    }
  } else {
    // ===================== 16 epilogue warps =====================
    const int quad = warp & 3;
    const int group = (warp - 2) >> 2;
    const int c = group * 32;                          // this warp's 32-column chunk of the tile
    const bool active = c < p.n_tile;
    const bool relu = p.relu != 0;
    const bool mask = p.aux_relu_mask != 0;
    const bool storer = (warp == 2 && lane == 0);
    const int r = quad * 32 + lane;                    // tile-local row
    // swizzled position of this thread's four 16-byte pieces inside its [128 x 128 B] block
    const uint32_t blk_off = uint32_t(c >> 6) * kChunkTileBytes + uint32_t(r) * 128;
    const uint32_t piece0 = uint32_t((c & 63) >> 3);
    int acc = 0;
    uint32_t acc_phase = 0;
    int dbg_tile = -1;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      ++dbg_tile;
      const int m_blk = t / p.num_n_tiles, n_blk = t % p.num_n_tiles;
      const bool row_ok = (m_blk * kBlockM + r) < p.M;
      const int n0 = n_blk * p.n_tile;
      uint8_t* cbuf = c_stage + acc * c_bytes;
      // the TMA store (and the statistics MMAs) that last read this staging buffer, two tiles ago,
      // must have drained it
      if (storer) { RDP_DBG_STAMP(5); }
      if (storer) tma_store_wait_read<1>();
      if (storer) { RDP_DBG_STAMP(6); }       // staging buffer drained by the store of two tiles ago
      if (kStats) mbar_wait(&stats_done[acc], acc_phase ^ 1);
      mbar_wait(&tmem_full[acc], acc_phase);
      if (kAux) mbar_wait(&x_full[acc], acc_phase);
      tc_fence_after();
      if (storer) { RDP_DBG_STAMP(7); }       // accumulator (and aux) ready
      epi_barrier();
      if (storer) { RDP_DBG_STAMP(8); }
      if (active) {
        const int col0 = n0 + c;
        uint32_t raw[32];
        tmem_ld_32x32(tmem_base + (uint32_t(quad * 32) << 16) + uint32_t(acc) * kFastNTile + uint32_t(c), raw);
        uint4 ua[4];
        if (kAux) {
          const uint8_t* xrow = x_stage + acc * c_bytes + blk_off;
#pragma unroll
          for (int q = 0; q < 4; ++q)
            ua[q] = *reinterpret_cast<const uint4*>(xrow + (((piece0 + q) ^ uint32_t(r & 7)) << 4));
        }
        tmem_ld_wait();
        float v[32];
        const float4* b4 = reinterpret_cast<const float4*>(s_bias + col0);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 bb = b4[q];
          v[q * 4 + 0] = __uint_as_float(raw[q * 4 + 0]) + bb.x;
          v[q * 4 + 1] = __uint_as_float(raw[q * 4 + 1]) + bb.y;
          v[q * 4 + 2] = __uint_as_float(raw[q * 4 + 2]) + bb.z;
          v[q * 4 + 3] = __uint_as_float(raw[q * 4 + 3]) + bb.w;
        }
        if (relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (kAux && bn_bwd) {
          // dgrad fused with the BatchNorm + ReLU backward of the layer it feeds: acc is dL/dy (grad wrt the
          // BN output), aux is r (the ReLU output = BN input); per-column alpha/beta/delta carry the batch
          // statistics of dy, which the caller derived from the layer's weight gradient (no pass over dy)
          const uint32_t* au = reinterpret_cast<const uint32_t*>(ua);
          const float4* al4 = reinterpret_cast<const float4*>(s_coef + col0);
          const float4* be4 = reinterpret_cast<const float4*>(s_coef + kBnCoefCols + col0);
          const float4* de4 = reinterpret_cast<const float4*>(s_coef + 2 * kBnCoefCols + col0);
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 al = al4[q], be = be4[q], de = de4[q];
            const float a0 = __uint_as_float(au[2 * q] << 16), a1 = __uint_as_float(au[2 * q] & 0xffff0000u);
            const float a2 = __uint_as_float(au[2 * q + 1] << 16), a3 = __uint_as_float(au[2 * q + 1] & 0xffff0000u);
            v[4 * q + 0] = a0 > 0.f ? fmaf(al.x, v[4 * q + 0], fmaf(be.x, a0, de.x)) : 0.f;
            v[4 * q + 1] = a1 > 0.f ? fmaf(al.y, v[4 * q + 1], fmaf(be.y, a1, de.y)) : 0.f;
            v[4 * q + 2] = a2 > 0.f ? fmaf(al.z, v[4 * q + 2], fmaf(be.z, a2, de.z)) : 0.f;
            v[4 * q + 3] = a3 > 0.f ? fmaf(al.w, v[4 * q + 3], fmaf(be.w, a3, de.w)) : 0.f;
          }
        } else if (kAux && mask) {
          const uint32_t* au = reinterpret_cast<const uint32_t*>(ua);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float a0 = __uint_as_float(au[i] << 16), a1 = __uint_as_float(au[i] & 0xffff0000u);
            v[2 * i] = a0 > 0.f ? v[2 * i] : 0.f;
            v[2 * i + 1] = a1 > 0.f ? v[2 * i + 1] : 0.f;
          }
        }
        uint32_t packed[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) packed[i] = row_ok ? pack_bf16x2(v[2 * i], v[2 * i + 1]) : 0u;
        uint8_t* crow = cbuf + blk_off;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4*>(crow + (((piece0 + q) ^ uint32_t(r & 7)) << 4)) =
              make_uint4(packed[q * 4], packed[q * 4 + 1], packed[q * 4 + 2], packed[q * 4 + 3]);
      }
      if (storer) { RDP_DBG_STAMP(9); }       // tile converted and staged (this warp)
      tc_fence_before();
      mbar_arrive(&tmem_empty[acc]);
      if (kAux && kStats != 2) mbar_arrive(&x_empty[acc]);
      fence_async_smem();                       // generic-proxy smem writes -> visible to TMA / UMMA
      if (storer) { RDP_DBG_STAMP(10); }
      epi_barrier();
      if (storer) { RDP_DBG_STAMP(11); }
      if (storer) {
        for (int ch = 0; ch < n_chunks64; ++ch)
          tma_store_2d(&tmap_c, cbuf + ch * kChunkTileBytes, n0 + ch * 64, m_blk * kBlockM);
        tma_store_commit();
        if (kStats) mbar_arrive(&c_ready[acc]);
        RDP_DBG_STAMP(12);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (storer) tma_store_wait_all();
    if (kStats && group == 0 && blockIdx.x < num_tiles) {
      // every tile of this CTA has the same n_blk (grid stride is a multiple of num_n_tiles), so the
      // TMEM statistics accumulators cover columns [n0, n0 + n_tile).  Lane l of quadrant q owns row
      // 32q + l: it reads the diagonal elements (32q + l, 32q + l) of D_s and D_g.
      mbar_wait(stats_final, 0);
      tc_fence_after();
      const int n0 = (blockIdx.x % p.num_n_tiles) * p.n_tile;
      if (quad * 32 < p.n_tile) {
        uint32_t ds[32], dg[32];
        const uint32_t lanes = uint32_t(quad * 32) << 16;
        tmem_ld_32x32(tmem_base + lanes + 2 * kFastNTile + uint32_t(quad * 32), ds);
        if (kStats != 3) tmem_ld_32x32(tmem_base + lanes + 3 * kFastNTile + uint32_t(quad * 32), dg);
        tmem_ld_wait();
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (lane == j) { s0 = __uint_as_float(ds[j]); s1 = (kStats != 3) ? __uint_as_float(dg[j]) : 0.f; }
        }
        const int col = n0 + quad * 32 + lane;
        atomicAdd(&p.stats[col], s0);
        if (kStats != 3 && p.stats1 != nullptr) atomicAdd(&p.stats1[col], s1);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kStats ? 512 : 256>(tmem_base);
  }
}

template <int kStats, bool kAux>
static int launch_fast(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
                       const GemmTnParams& p, cudaStream_t stream) {
  RDP_SMEM_ATTR_ONCE_PER_DEVICE((gemm_tn_fast_kernel<kStats, kAux>), 227 * 1024);
  const size_t smem = size_t(kAux ? 2 : 3) * (16384 + kFastNTile * 128) + (kAux ? 4 : 2) * 2 * kChunkTileBytes +
                      (kStats ? kChunkTileBytes : 0) + 256 + kStatsCols * sizeof(float) +
                      (kAux ? 3 * kBnCoefCols * sizeof(float) : 0) + 1024;
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  // a CTA must only ever see one n_blk (its TMEM statistics accumulators are per column block)
  int grid = (num_sms() / p.num_n_tiles) * p.num_n_tiles;
  if (grid < p.num_n_tiles) grid = p.num_n_tiles;
  if (tiles < grid) grid = tiles;
  static long long* dbg_buf = nullptr;
  static int dbg_left = -1;
  if (dbg_left < 0) dbg_left = getenv("RDP_GEMM_DBG") ? atoi(getenv("RDP_GEMM_DBG")) : 0;
  GemmTnParams q = p;
  q.dbg = nullptr;
  if (dbg_left > 0) {
    if (!dbg_buf) RDP_CHECK_CUDA(cudaMalloc(&dbg_buf, 32 * 16 * sizeof(long long)));
    RDP_CHECK_CUDA(cudaMemsetAsync(dbg_buf, 0, 32 * 16 * sizeof(long long), stream));
    q.dbg = dbg_buf;
  }
  RDP_CHECK_CUDA(launch_pdl(gemm_tn_fast_kernel<kStats, kAux>, dim3(grid), dim3(kFastThreads), smem, stream, ta, tb,
                            tc, tx, q));
  RDP_CHECK_CUDA(cudaGetLastError());
  if (q.dbg) {
    --dbg_left;
    static long long host[32 * 16];
    RDP_CHECK_CUDA(cudaStreamSynchronize(stream));
    RDP_CHECK_CUDA(cudaMemcpy(host, dbg_buf, sizeof(host), cudaMemcpyDeviceToHost));
    long long t0 = host[0];
    fprintf(stderr, "[rdp gemm dbg] stats=%d aux=%d M=%d N=%d K=%d grid=%d (ns since the producer's first stamp)\n",
            kStats, int(kAux), p.M, p.N, p.K, grid);
    for (int t = 0; t < 32 && host[t * 16] != 0; ++t) {
      fprintf(stderr, "  tile %2d:", t);
      for (int k = 0; k < 13; ++k) fprintf(stderr, " %6lld", host[t * 16 + k] ? host[t * 16 + k] - t0 : -1);
      fprintf(stderr, "\n");
    }
  }
  return 0;
}

// ----------------------------------------------------------------------------------------------
// wgrad: D[Mw, Nw] += sum over batch rows of G[r, :Mw]^T X[r, :Nw]   (both MN-major)
// ----------------------------------------------------------------------------------------------
struct WgradParams {
  int rows;            // batch rows (reduction length)
  int Mw, Nw;          // output extents (Mw <= 128 per CTA tile, Nw <= 256 per CTA tile)
  int n_tile;          // multiple of 16, <= 256
  int rows_per_split;  // multiple of 64
  float* out;          // [Mw_total, ldo] fp32, accumulated with atomics
  int ldo;
  float* colsum;       // optional [Mw_total]: sum over rows of G (bias gradient), or null
  // deterministic mode: every split STORES its tile at partial + split * partial_stride (same [m, ldo] layout as
  // `out`); wgrad_reduce_kernel then adds the splits in order.  With one split the tile is stored into `out`.
  float* partial;
  long long partial_stride;
};

constexpr int kWgradThreads = 192;
constexpr int kWgradStages = 4;

__global__ void __launch_bounds__(kWgradThreads, 1)
gemm_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_g, const __grid_constant__ CUtensorMap tmap_x,
                  const WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int n_chunks = (p.n_tile + 63) / 64;                 // 64-wide MN chunks of the X operand
  const uint32_t chunk_bytes = kBlockK * 128;                // 64 k-rows x 128 B = 8 KB
  const uint32_t a_bytes = 2 * chunk_bytes;                  // 128 output rows = 2 chunks of G
  const uint32_t b_bytes = uint32_t(n_chunks) * chunk_bytes;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  uint8_t* tiles = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + kWgradStages * stage_bytes);
  uint64_t* empty_bar = full_bar + kWgradStages;
  uint64_t* tmem_full = empty_bar + kWgradStages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_trigger();
  const int split = blockIdx.x;
  const int m_blk = blockIdx.y;
  const int n_blk = blockIdx.z;
  const int r_begin = split * p.rows_per_split;
  const int r_end = min(p.rows, r_begin + p.rows_per_split);
  const int num_kb = (r_end > r_begin) ? (r_end - r_begin + kBlockK - 1) / kBlockK : 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_g);
    tma_prefetch_desc(&tmap_x);
    for (int s = 0; s < kWgradStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                         // operands and the gradient buffer belong to earlier kernels
  const uint32_t tmem_base = *tmem_ptr;

  if (num_kb > 0) {
    if (warp == 0) {
      if (lane == 0) {
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = tiles + stage * stage_bytes;
          uint8_t* b_dst = a_dst + a_bytes;
          mbar_expect_tx(&full_bar[stage], stage_bytes);
          const int r = r_begin + kb * kBlockK;
          // rows past r_end inside the last k-block belong to the next split: they are loaded
          // by that split too, so clamp by giving this split whole k-blocks only (host side
          // guarantees rows_per_split % 64 == 0, hence no overlap); rows >= p.rows read as 0.
          tma_load_2d(a_dst, &tmap_g, &full_bar[stage], m_blk * 128, r);
          tma_load_2d(a_dst + chunk_bytes, &tmap_g, &full_bar[stage], m_blk * 128 + 64, r);
          for (int ch = 0; ch < n_chunks; ++ch)
            tma_load_2d(b_dst + ch * chunk_bytes, &tmap_x, &full_bar[stage], n_blk * p.n_tile + ch * 64, r);
          if (++stage == kWgradStages) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        const uint32_t idesc = umma_idesc_bf16(128, uint32_t(p.n_tile), true, true);
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(tiles + stage * stage_bytes);
          const uint32_t b_addr = a_addr + a_bytes;
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // 16 reduction rows = 2048 B further down the tile; chunks are LBO = 8 KB apart
            const uint64_t a_desc = umma_smem_desc(a_addr + k * 2048, chunk_bytes, 1024);
            const uint64_t b_desc = umma_smem_desc(b_addr + k * 2048, chunk_bytes, 1024);
            umma_bf16(tmem_base, a_desc, b_desc, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == num_kb - 1) umma_commit(tmem_full);
          if (++stage == kWgradStages) { stage = 0; phase ^= 1; }
        }
      }
    } else {
      const int quad = warp & 3;
      const int m = m_blk * 128 + quad * 32 + lane;
      mbar_wait(tmem_full, 0);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (uint32_t(quad * 32) << 16);
      for (int c = 0; c < p.n_tile; c += 32) {
        const int col0 = n_blk * p.n_tile + c;
        if (col0 >= p.Nw) break;
        uint32_t raw[32];
        tmem_ld_32x32(t_row + uint32_t(c), raw);
        tmem_ld_wait();
        if (m < p.Mw && p.partial != nullptr) {
          float* dst = p.partial + size_t(split) * p.partial_stride + size_t(m) * p.ldo + col0;
          const int ncols = min(32, min(p.n_tile - c, p.Nw - col0));
          if ((p.ldo & 3) == 0 && ncols == 32) {
#pragma unroll
            for (int q = 0; q < 8; ++q)
              *reinterpret_cast<uint4*>(dst + q * 4) = make_uint4(raw[q * 4], raw[q * 4 + 1], raw[q * 4 + 2], raw[q * 4 + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < ncols) dst[j] = __uint_as_float(raw[j]);
          }
        } else if (m < p.Mw) {
          float* dst = p.out + size_t(m) * p.ldo + col0;
          const int ncols = min(32, min(p.n_tile - c, p.Nw - col0));
          if ((p.ldo & 3) == 0) {
            // red.global.add.v4.f32: a quarter of the reduction instructions of scalar atomics
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              if (q * 4 + 3 < ncols) {
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + q * 4),
                             "f"(__uint_as_float(raw[q * 4])), "f"(__uint_as_float(raw[q * 4 + 1])),
                             "f"(__uint_as_float(raw[q * 4 + 2])), "f"(__uint_as_float(raw[q * 4 + 3]))
                             : "memory");
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (q * 4 + j < ncols) atomicAdd(dst + q * 4 + j, __uint_as_float(raw[q * 4 + j]));
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < ncols) atomicAdd(dst + j, __uint_as_float(raw[j]));
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

// out[i] = partial[0][i] + partial[1][i] + ... in split order: the same bits whatever the schedule was.
// Several weight matrices (all layers of a step) per launch; four split loads in flight per thread.
constexpr int kReduceMaxSeg = 16;
struct WgradReduceParams {
  const float* partial[kReduceMaxSeg];
  float* out[kReduceMaxSeg];
  long long elems[kReduceMaxSeg];     // = stride between splits
  int splits[kReduceMaxSeg];
  int first_block[kReduceMaxSeg + 1];
  int n_seg;
};

__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const __grid_constant__ WgradReduceParams p) {
  int seg = 0;
  while (seg + 1 < p.n_seg && int(blockIdx.x) >= p.first_block[seg + 1]) ++seg;
  const float* __restrict__ partial = p.partial[seg];
  float* __restrict__ out = p.out[seg];
  const long long n = p.elems[seg], stride = n;
  const int splits = p.splits[seg];
  const int nblk = p.first_block[seg + 1] - p.first_block[seg];
  const long long tid = (blockIdx.x - p.first_block[seg]) * (long long)blockDim.x + threadIdx.x;
  const long long nthreads = (long long)nblk * blockDim.x;
  const bool vec = (n & 3) == 0 && (reinterpret_cast<uintptr_t>(partial) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  const long long n4 = vec ? n / 4 : 0;
  for (long long i = tid; i < n4; i += nthreads) {
    float4 acc = *reinterpret_cast<const float4*>(partial + i * 4);
    int s = 1;
    for (; s + 3 < splits; s += 4) {
      const float4 v0 = *reinterpret_cast<const float4*>(partial + (s + 0) * stride + i * 4);
      const float4 v1 = *reinterpret_cast<const float4*>(partial + (s + 1) * stride + i * 4);
      const float4 v2 = *reinterpret_cast<const float4*>(partial + (s + 2) * stride + i * 4);
      const float4 v3 = *reinterpret_cast<const float4*>(partial + (s + 3) * stride + i * 4);
      acc.x += v0.x; acc.y += v0.y; acc.z += v0.z; acc.w += v0.w;
      acc.x += v1.x; acc.y += v1.y; acc.z += v1.z; acc.w += v1.w;
      acc.x += v2.x; acc.y += v2.y; acc.z += v2.z; acc.w += v2.w;
      acc.x += v3.x; acc.y += v3.y; acc.z += v3.z; acc.w += v3.w;
    }
    for (; s < splits; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(partial + s * stride + i * 4);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    *reinterpret_cast<float4*>(out + i * 4) = acc;
  }
  for (long long i = n4 * 4 + tid; i < n; i += nthreads) {   // scalar route (unaligned or odd sizes)
    float acc = partial[i];
    for (int s = 1; s < splits; ++s) acc += partial[s * stride + i];
    out[i] = acc;
  }
}

static int launch_wgrad_reduce(const float* const* ws, const long long* elems, const int* splits, float* const* outs,
                               int n, cudaStream_t stream) {
  WgradReduceParams p;
  memset(&p, 0, sizeof(p));
  int blocks = 0, seg = 0;
  for (int i = 0; i < n; ++i) {
    if (splits[i] <= 1) continue;                 // a single split was stored straight into its output
    p.partial[seg] = ws[i]; p.out[seg] = outs[i]; p.elems[seg] = elems[i]; p.splits[seg] = splits[i];
    p.first_block[seg] = blocks;
    long long b = (elems[i] / 4 + 255) / 256;
    if (b < 1) b = 1;
    if (b > 1024) b = 1024;
    blocks += int(b);
    ++seg;
  }
  if (seg == 0) return 0;
  p.first_block[seg] = blocks;
  p.n_seg = seg;
  wgrad_reduce_kernel<<<blocks, 256, 0, stream>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static size_t gemm_tn_smem_bytes(int n_tile) {
  const size_t a_bytes = kBlockM * kBlockK * 2;
  const size_t b_bytes = (size_t(n_tile) * kBlockK * 2 + 1023) & ~size_t(1023);
  return kStages * (a_bytes + b_bytes) + 256 + 3 * kStatsCols * sizeof(float) + 1024;
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_gemm_tn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int out_f32,
                int M, int N, int K, const float* bias, int relu, const void* aux, int ld_aux,
                int aux_relu_mask, int stats_mode, float* stats, void* stream) {
  RDP_REQUIRE(M > 0 && N > 0 && K > 0, "rdp_gemm_tn: empty problem %dx%dx%d", M, N, K);
  RDP_REQUIRE(N % 16 == 0, "rdp_gemm_tn: N=%d must be a multiple of 16", N);
  RDP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "rdp_gemm_tn: pitches must be multiples of 8");
  RDP_REQUIRE(stats_mode == 0 || stats != nullptr, "rdp_gemm_tn: stats buffer missing");
  RDP_REQUIRE(!(stats_mode == 2 || aux_relu_mask) || (aux != nullptr && ld_aux % 8 == 0),
              "rdp_gemm_tn: aux matrix missing or mis-pitched");
  GemmTnParams p;
  p.M = M; p.N = N; p.K = K;
  p.n_tile = N >= kMaxNTile ? kMaxNTile : N;
  if (N > kMaxNTile && N % kMaxNTile != 0) p.n_tile = (N % 128 == 0) ? 128 : ((N % 64 == 0) ? 64 : 16);
  p.num_m_tiles = (M + kBlockM - 1) / kBlockM;
  p.num_n_tiles = (N + p.n_tile - 1) / p.n_tile;
  p.num_k_blocks = (K + kBlockK - 1) / kBlockK;
  p.bias = bias; p.relu = relu; p.out = out; p.out_f32 = out_f32; p.ldc = ldc;
  p.aux = static_cast<const __nv_bfloat16*>(aux); p.ld_aux = ld_aux; p.aux_relu_mask = aux_relu_mask;
  p.stats_mode = stats_mode; p.stats = stats;
  p.stats1 = stats ? stats + N : nullptr; p.bn_alpha = p.bn_beta = p.bn_delta = nullptr; p.dbg = nullptr;
  CUtensorMap ta, tb;
  int rc = make_tmap_2d_bf16(&ta, a, M, K, lda, kBlockM, kBlockK);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tb, b, N, K, ldb, p.n_tile, kBlockK);
  if (rc) return rc;
  const size_t smem = gemm_tn_smem_bytes(p.n_tile);
  RDP_SMEM_ATTR_ONCE_PER_DEVICE(gemm_tn_kernel, 227 * 1024);
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  const bool use_aux = (stats_mode == 2) || aux_relu_mask;
  // large plain GEMMs: CTA-pair kernel, 256 x 256 tiles (at least half a wave of clusters, deep K)
  if (!out_f32 && stats_mode == 0 && !use_aux && N % 256 == 0 && K >= 512 &&
      ((M + 255) / 256) * (N / 256) >= num_sms() / 4)
    return gemm2_tn(a, lda, b, ldb, out, ldc, M, N, K, bias, relu, 0, cs);
  if (!out_f32 && N % 64 == 0 && N <= kStatsCols) {
    GemmTnParams f = p;
    f.n_tile = (N % kFastNTile == 0) ? kFastNTile : 64;
    f.num_n_tiles = N / f.n_tile;
    CUtensorMap fb, fc, fx;
    rc = make_tmap_2d_bf16(&fb, b, N, K, ldb, f.n_tile, kBlockK);
    if (rc) return rc;
    rc = make_tmap_2d_bf16(&fc, out, M, N, ldc, kBlockM, 64);
    if (rc) return rc;
    fx = fc;
    if (use_aux) {
      rc = make_tmap_2d_bf16(&fx, aux, M, N, ld_aux, kBlockM, 64);
      if (rc) return rc;
    }
    if (stats_mode == 0 && !use_aux) return launch_fast<0, false>(ta, fb, fc, fx, f, cs);
    if (stats_mode == 1 && !use_aux) return launch_fast<1, false>(ta, fb, fc, fx, f, cs);
    if (stats_mode == 2) return launch_fast<2, true>(ta, fb, fc, fx, f, cs);
    if (stats_mode == 0 && use_aux) return launch_fast<0, true>(ta, fb, fc, fx, f, cs);
    if (stats_mode == 1 && use_aux) return launch_fast<1, true>(ta, fb, fc, fx, f, cs);   // ReLU-masked dgrad + column sums
  }
  gemm_tn_kernel<<<grid, kGemmThreads, smem, cs>>>(ta, tb, p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_gemm_tn_bnbwd(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K,
                      const void* aux, int ld_aux, const float* alpha, const float* beta, const float* delta,
                      float* colsum, void* stream) {
  RDP_REQUIRE(M > 0 && N > 0 && K > 0, "rdp_gemm_tn_bnbwd: empty problem %dx%dx%d", M, N, K);
  RDP_REQUIRE(N % 64 == 0 && N <= kBnCoefCols, "rdp_gemm_tn_bnbwd: N=%d must be a multiple of 64 and <= %d", N,
              kBnCoefCols);
  RDP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0 && ld_aux % 8 == 0,
              "rdp_gemm_tn_bnbwd: pitches must be multiples of 8");
  RDP_REQUIRE(aux && alpha && beta && delta && colsum, "rdp_gemm_tn_bnbwd: missing operand");
  GemmTnParams f;
  f.M = M; f.N = N; f.K = K;
  f.n_tile = (N % kFastNTile == 0) ? kFastNTile : 64;
  f.num_m_tiles = (M + kBlockM - 1) / kBlockM;
  f.num_n_tiles = N / f.n_tile;
  f.num_k_blocks = (K + kBlockK - 1) / kBlockK;
  f.bias = nullptr; f.relu = 0; f.out = out; f.out_f32 = 0; f.ldc = ldc;
  f.aux = static_cast<const __nv_bfloat16*>(aux); f.ld_aux = ld_aux; f.aux_relu_mask = 0;
  f.stats_mode = 1; f.stats = colsum; f.stats1 = nullptr;
  f.bn_alpha = alpha; f.bn_beta = beta; f.bn_delta = delta; f.dbg = nullptr;
  CUtensorMap ta, fb, fc, fx;
  int rc = make_tmap_2d_bf16(&ta, a, M, K, lda, kBlockM, kBlockK);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&fb, b, N, K, ldb, f.n_tile, kBlockK);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&fc, out, M, N, ldc, kBlockM, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&fx, aux, M, N, ld_aux, kBlockM, 64);
  if (rc) return rc;
  return launch_fast<3, true>(ta, fb, fc, fx, f, static_cast<cudaStream_t>(stream));    // column sums only
}

int rdp_gemm_wgrad(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows,
                   int Mw, int Nw, void* stream) {
  RDP_REQUIRE(rows > 0 && Mw > 0 && Nw > 0, "rdp_gemm_wgrad: empty problem");
  RDP_REQUIRE(ldg % 8 == 0 && ldx % 8 == 0, "rdp_gemm_wgrad: pitches must be multiples of 8");
  if (ldo % 4 == 0 && Nw % 4 == 0 && rows >= 512 && ((Mw + 255) / 256) * ((Nw + 255) / 256) >= num_sms() / 4)
    return gemm2_wgrad(g, ldg, x, ldx, out, ldo, rows, Mw, Nw, 0, static_cast<cudaStream_t>(stream));
  WgradParams p;
  p.rows = rows; p.Mw = Mw; p.Nw = Nw;
  int n_tile = ((Nw + 15) / 16) * 16;
  if (n_tile > 256) n_tile = 256;
  p.n_tile = n_tile;
  const int m_tiles = (Mw + 127) / 128;
  const int n_tiles = (Nw + n_tile - 1) / n_tile;
  int splits = num_sms() / (m_tiles * n_tiles);
  if (splits < 1) splits = 1;
  int rps = (rows + splits - 1) / splits;
  rps = ((rps + kBlockK - 1) / kBlockK) * kBlockK;
  if (rps < 4 * kBlockK) rps = 4 * kBlockK;
  splits = (rows + rps - 1) / rps;
  p.rows_per_split = rps;
  p.out = out; p.ldo = ldo; p.colsum = nullptr; p.partial = nullptr; p.partial_stride = 0;
  CUtensorMap tg, tx;
  int rc = make_tmap_2d_bf16(&tg, g, rows, Mw, ldg, kBlockK, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tx, x, rows, Nw, ldx, kBlockK, 64);
  if (rc) return rc;
  const int n_chunks = (n_tile + 63) / 64;
  const size_t smem = size_t(kWgradStages) * (2 + n_chunks) * (kBlockK * 128) + 256 + 1024;
  RDP_SMEM_ATTR_ONCE_PER_DEVICE(gemm_wgrad_kernel, 227 * 1024);
  dim3 grid(splits, m_tiles, n_tiles);
  RDP_CHECK_CUDA(launch_pdl(gemm_wgrad_kernel, grid, dim3(kWgradThreads), smem, static_cast<cudaStream_t>(stream), tg,
                            tx, p));
  return 0;
}

// Deterministic wgrad: out[Mw, Nw] (contiguous, ldo == Nw) = G^T X, bit-identical run to run.
// The batch is split over CTAs as above, but each split stores its tile into `ws` and the splits are added
// in order; the number of splits is limited by ws_floats (>= Mw * Nw: one split, stored straight into out).
static int wgrad_det_impl(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw,
                          int Nw, float* ws, long long ws_floats, int* splits_out, bool reduce_now, void* stream) {
  RDP_REQUIRE(rows > 0 && Mw > 0 && Nw > 0, "rdp_gemm_wgrad_det: empty problem");
  RDP_REQUIRE(ldg % 8 == 0 && ldx % 8 == 0, "rdp_gemm_wgrad_det: pitches must be multiples of 8");
  RDP_REQUIRE(ldo == Nw, "rdp_gemm_wgrad_det: out must be contiguous (ldo=%d Nw=%d)", ldo, Nw);
  WgradParams p;
  p.rows = rows; p.Mw = Mw; p.Nw = Nw;
  int n_tile = ((Nw + 15) / 16) * 16;
  if (n_tile > 256) n_tile = 256;
  p.n_tile = n_tile;
  const int m_tiles = (Mw + 127) / 128;
  const int n_tiles = (Nw + n_tile - 1) / n_tile;
  const long long elems = (long long)Mw * Nw;
  int splits = num_sms() / (m_tiles * n_tiles);
  const long long cap = ws != nullptr ? ws_floats / elems : 1;
  if (splits > cap) splits = int(cap);
  if (splits < 1) splits = 1;
  int rps = (rows + splits - 1) / splits;
  rps = ((rps + kBlockK - 1) / kBlockK) * kBlockK;
  if (rps < 4 * kBlockK) rps = 4 * kBlockK;
  splits = (rows + rps - 1) / rps;
  p.rows_per_split = rps;
  p.out = out; p.ldo = ldo; p.colsum = nullptr;
  p.partial = splits == 1 ? out : ws;
  p.partial_stride = elems;
  CUtensorMap tg, tx;
  int rc = make_tmap_2d_bf16(&tg, g, rows, Mw, ldg, kBlockK, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tx, x, rows, Nw, ldx, kBlockK, 64);
  if (rc) return rc;
  const int n_chunks = (n_tile + 63) / 64;
  const size_t smem = size_t(kWgradStages) * (2 + n_chunks) * (kBlockK * 128) + 256 + 1024;
  RDP_SMEM_ATTR_ONCE_PER_DEVICE(gemm_wgrad_kernel, 227 * 1024);
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  dim3 grid(splits, m_tiles, n_tiles);
  RDP_CHECK_CUDA(launch_pdl(gemm_wgrad_kernel, grid, dim3(kWgradThreads), smem, cs, tg, tx, p));
  if (splits_out != nullptr) *splits_out = splits;
  if (reduce_now && splits > 1) {
    const float* wsp = ws;
    return launch_wgrad_reduce(&wsp, &elems, &splits, &out, 1, cs);
  }
  return 0;
}

int rdp_gemm_wgrad_det(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw,
                       int Nw, float* ws, long long ws_floats, void* stream) {
  return wgrad_det_impl(g, ldg, x, ldx, out, ldo, rows, Mw, Nw, ws, ws_floats, nullptr, true, stream);
}

// Two-phase form for a whole backward pass: every layer stores its split tiles into its own scratch
// (rdp_gemm_wgrad_partial; *splits_out = number of splits, 1 = already stored into out), then ONE launch adds the
// splits of all layers in order (rdp_wgrad_reduce_many; up to 16 layers per launch).
int rdp_gemm_wgrad_partial(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw,
                           int Nw, float* ws, long long ws_floats, int* splits_out, void* stream) {
  return wgrad_det_impl(g, ldg, x, ldx, out, ldo, rows, Mw, Nw, ws, ws_floats, splits_out, false, stream);
}

int rdp_wgrad_reduce_many(const float* const* ws, const long long* elems, const int* splits, float* const* outs,
                          int n, void* stream) {
  RDP_REQUIRE(n >= 0 && n <= kReduceMaxSeg, "rdp_wgrad_reduce_many: at most %d matrices per call (got %d)",
              kReduceMaxSeg, n);
  return launch_wgrad_reduce(ws, elems, splits, outs, n, static_cast<cudaStream_t>(stream));
}

}  // extern "C"

// ================================================================================================
// Allocation-free entry point
// ================================================================================================
// rdp_gemm_tn needs a caller-provided [2, N] fp32 scratch for its per-column batch statistics, and
// the caller has to allocate and zero it.  Some callers cannot allocate device memory at all: a
// plugin living inside a host application whose device memory is owned end-to-end by that
// application's own pool (a stray cudaMalloc there can fail, fragment the pool or break its
// accounting), and a latency-critical loop that must never hit cudaMalloc's implicit device-wide
// synchronisation.  For them the scratch lives in the library image itself: a file-scope
// __device__ array that every GPU loading this library carries until the process exits
// (kNoAllocMaxN columns -> 8 KB per GPU).
//
// rdp_gemm_tn_bnstats_noalloc computes  out = relu?(a . b^T + bias)  exactly like rdp_gemm_tn and
// returns the per-column batch mean and biased variance of `out` in caller-provided outputs.
// It calls no cudaMalloc / cudaFree and no allocating library routine.
//
//  * Limit: N <= kNoAllocMaxN (1024); larger N is rejected with an error, never truncated.
//  * Set-up: the first call on each device resolves the device address of the scratch array with
//    cudaGetSymbolAddress (the array's name in host code is only a host-side stand-in, not an
//    address a kernel may use).  Make that first call per device outside any stream capture.
//  * Every call shares the one array.  It is cleared at the start of each call, on the caller's
//    stream, so back-to-back calls on one stream are safe; two calls that overlap on the GPU
//    (different streams, or different host threads on different streams) race on it.
namespace rdp {

constexpr int kNoAllocMaxN = 1024;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_noalloc_stats[2 * kNoAllocMaxN];

static float* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};
static std::mutex g_noalloc_mutex;

// mean[j] = sum / rows ; var[j] = max(sumsq / rows - mean^2, 0)   (device code may name the array)
__global__ void noalloc_finalize_kernel(const float* __restrict__ stats, int n, float rows,
                                        float* __restrict__ mean, float* __restrict__ var) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const float m = stats[j] / rows;
  mean[j] = m;
  var[j] = fmaxf(stats[n + j] / rows - m * m, 0.f);
}

}  // namespace rdp

extern "C" int rdp_gemm_tn_bnstats_noalloc(const void* a, int lda, const void* b, int ldb, void* out, int ldc,
                                           int out_f32, int M, int N, int K, const float* bias, int relu,
                                           float* mean_out, float* var_out, void* stream) {
  RDP_REQUIRE(M > 0 && N > 0 && K > 0, "rdp_gemm_tn_bnstats_noalloc: empty problem %dx%dx%d", M, N, K);
  RDP_REQUIRE(N <= kNoAllocMaxN,
              "rdp_gemm_tn_bnstats_noalloc: N=%d exceeds the built-in scratch (max %d columns); use "
              "rdp_gemm_tn with your own [2, N] statistics buffer", N, kNoAllocMaxN);
  RDP_REQUIRE(N % 16 == 0, "rdp_gemm_tn_bnstats_noalloc: N=%d must be a multiple of 16", N);
  RDP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0,
              "rdp_gemm_tn_bnstats_noalloc: pitches must be multiples of 8");
  RDP_REQUIRE(mean_out != nullptr && var_out != nullptr, "rdp_gemm_tn_bnstats_noalloc: outputs missing");
  int dev = 0;
  RDP_CHECK_CUDA(cudaGetDevice(&dev));
  RDP_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "rdp_gemm_tn_bnstats_noalloc: device %d out of range", dev);
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
      // once per device, with that device current: host code must not pass the symbol itself
      float* ptr = nullptr;
      RDP_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_stats));
      RDP_CHECK_CUDA(cudaFuncSetAttribute(gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
      g_noalloc_ptr[dev] = ptr;
    }
    scratch = g_noalloc_ptr[dev];
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the array keeps the previous call's sums: clear the part this call accumulates into
  RDP_CHECK_CUDA(cudaMemsetAsync(scratch, 0, size_t(2) * N * sizeof(float), s));
  GemmTnParams p;
  p.M = M; p.N = N; p.K = K;
  p.n_tile = N >= kMaxNTile ? kMaxNTile : N;
  if (N > kMaxNTile && N % kMaxNTile != 0) p.n_tile = (N % 128 == 0) ? 128 : ((N % 64 == 0) ? 64 : 16);
  p.num_m_tiles = (M + kBlockM - 1) / kBlockM;
  p.num_n_tiles = (N + p.n_tile - 1) / p.n_tile;
  p.num_k_blocks = (K + kBlockK - 1) / kBlockK;
  p.bias = bias; p.relu = relu; p.out = out; p.out_f32 = out_f32; p.ldc = ldc;
  p.aux = nullptr; p.ld_aux = 0; p.aux_relu_mask = 0;
  p.stats_mode = 1; p.stats = scratch;
  p.stats1 = scratch + N; p.bn_alpha = p.bn_beta = p.bn_delta = nullptr; p.dbg = nullptr;
  CUtensorMap ta, tb;
  int rc = make_tmap_2d_bf16(&ta, a, M, K, lda, kBlockM, kBlockK);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tb, b, N, K, ldb, p.n_tile, kBlockK);
  if (rc) return rc;
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  gemm_tn_kernel<<<grid, kGemmThreads, gemm_tn_smem_bytes(p.n_tile), s>>>(ta, tb, p);
  RDP_CHECK_CUDA(cudaGetLastError());
  noalloc_finalize_kernel<<<(N + 255) / 256, 256, 0, s>>>(scratch, N, float(M), mean_out, var_out);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
