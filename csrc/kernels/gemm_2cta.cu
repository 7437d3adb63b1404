// Large-GEMM path: CTA-pair (cta_group::2) tcgen05 kernels.
//
// The 1-CTA kernels in gemm_tcgen05.cu move 32-48 KB of operands per 128 x {128,256} x 64 MMA
// block; at BigMLP sizes (8192 x 16384 x 8192) that makes them L2->SM bandwidth bound at about half
// of the tensor-core rate.  Here two CTAs of one cluster (one TPC) cooperate on UMMAs of
// M = 256, N = 256: each CTA stages its own rows of A and *half* of B (128 of the 256 columns),
// the pair reads both halves, and each CTA's TMEM receives its slice of the fp32 accumulator.
//
//   warp 0      TMA producer (both CTAs; completion bytes are signalled on the LEADER's barrier)
//   warp 1      MMA issuer (leader CTA) / tile scheduler (CTA 1, dynamic schedule); TMEM allocation
//               in both CTAs (cta_group::2)
//   warps 2..   epilogue: tcgen05.ld of the CTA's own accumulator slice (4 warps per 128-row half)
//
// Template parameters of gemm_pair_kernel:
//   kMode  0  out[M, N] (bf16) = act(A[M, K] . B[N, K]^T + bias)     K-major A and B (forward)
//          1  out[Mw, Nw] (fp32) += G[rows, Mw]^T . X[rows, Nw]      MN-major A and B (wgrad),
//             epilogue red.global.add.v4.f32 straight into weight.grad
//          2  out[M, N] (bf16) = A[M, K] . B[K, N]                   K-major A, MN-major B (dgrad from
//             the row-major weight, no transposed copy)
//   kMH    1  cluster tile 256 x 256, 6-stage TMA ring, TMEM double-buffered (epilogue hidden)
//          2  cluster tile 512 x 256, 4-stage ring, all 512 TMEM columns hold one tile; the epilogue
//             drains TMEM into registers first so its stores overlap the next tile's MMAs
//   kDynamic  tiles by cluster-launch-control work stealing instead of the static persistent loop
//
// Tiles are visited in groups of `group_m` row-blocks (column-major inside a group) so that the
// clusters in flight share a small set of A/B panels in the 126 MB L2.
#include "common.cuh"

#include <stdlib.h>

namespace rdp {

constexpr int k2Tile = 256;                          // cluster tile edge (M and N)
constexpr uint32_t k2HalfBytes = 128 * 64 * 2;       // one [128 x 64] bf16 operand block, 16 KB
constexpr uint32_t k2StoreBytes = 128 * 128;         // [128 rows x 64 cols] bf16 staging block

struct Gemm2Params {
  int M, N, K;                       // wgrad: M = Mw, N = Nw, K = batch rows
  int num_m_tiles, num_n_tiles, num_k_blocks;
  int group_m;
  const float* bias;
  int relu;
  float* out_f32;                    // wgrad only
  int ldo;
};

// ---- cluster / pair PTX ----------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_addr` (a shared::cta address) inside CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load issued by either CTA of the pair; `bar_cluster_addr` names the leader's mbarrier
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, uint32_t bar_cluster_addr,
                                                 int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_result) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(smem_result))
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at the same smem offset in BOTH CTAs once all prior MMAs have retired
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(uint16_t(3))
      : "memory");
}
__device__ __forceinline__ void tma_store_2d_b(const CUtensorMap* map, const void* smem_src, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}

__device__ __forceinline__ void tma_store_wait_read_le1() {
  asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
}

// ---- cluster launch control: hardware work stealing for persistent clusters -------------------
// The grid has one cluster per output tile.  A running cluster asks the hardware to cancel a cluster
// that has not been launched yet and, if that succeeds, computes the cancelled cluster's tile
// itself.  Unlike a static "tile += gridDim" loop this keeps working when only some clusters are
// resident -- e.g. while blocks of an overlapped all-reduce kernel occupy a few SMs: the resident
// clusters simply take more tiles, and the others join as SMs free up.
// One slot: a cluster claims at most ONE tile ahead of the one it is computing (its producer warp
// prefetches that tile's operands); deeper rings only hoard tiles and unbalance the tail.
constexpr int kClcSlots = 1;
__device__ __forceinline__ void clc_try_cancel(void* response16, uint64_t* bar) {
  asm volatile(
      "clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.multicast::cluster::all.b128"
      " [%0], [%1];" ::"r"(smem_u32(response16)), "r"(smem_u32(bar)) : "memory");
}
// tile index of the cancelled cluster (its first CTA id / 2), or -1 when nothing was left to cancel
__device__ __forceinline__ int clc_decode(const void* response16) {
  uint32_t x, valid;
  asm volatile(
      "{\n\t.reg .pred p1;\n\t.reg .b128 r;\n\t"
      "ld.shared.b128 r, [%2];\n\t"
      "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p1, r;\n\t"
      "selp.u32 %1, 1, 0, p1;\n\t"
      "mov.u32 %0, 0;\n\t"
      "@p1 clusterlaunchcontrol.query_cancel.get_first_ctaid::x.b32.b128 %0, r;\n\t}"
      : "=r"(x), "=r"(valid) : "r"(smem_u32(response16)) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  return valid ? int(x >> 1) : -1;
}
__device__ __forceinline__ void mbar_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.release.cluster.shared::cluster.b64 _, [%0], %1;"
               ::"r"(cluster_addr), "r"(bytes) : "memory");
}

// Per-role cursor over the tiles of this cluster: static stride, or the CLC response ring.
struct TileCursor {
  uint8_t* resp;          // [kClcSlots][16]   (this CTA's copy of the responses)
  uint64_t* full;         // [kClcSlots]
  uint64_t* empty;        // [kClcSlots]       (the SCHEDULER CTA's barriers are the ones that count)
  int slot;
  uint32_t phase;
  int stride, limit;      // static mode
  __device__ __forceinline__ int next(bool dynamic, int t, bool arriving_lane) {
    if (!dynamic) return (t + stride < limit) ? t + stride : -1;
    mbar_wait(&full[slot], phase);
    const int nt = clc_decode(resp + slot * 16);
    if (arriving_lane) mbar_arrive_cluster(map_to_cta(smem_u32(&empty[slot]), 1));
    if (++slot == kClcSlots) { slot = 0; phase ^= 1; }
    return nt;
  }
};

__device__ __forceinline__ void tile_coords(const Gemm2Params& p, int t, int& m_blk, int& n_blk) {
  const int per_group = p.group_m * p.num_n_tiles;
  const int g = t / per_group, in_g = t - g * per_group;
  const int first_m = g * p.group_m;
  const int gsz = min(p.group_m, p.num_m_tiles - first_m);
  m_blk = first_m + in_g % gsz;
  n_blk = in_g / gsz;
}

// kMH = number of 128-row accumulator halves per CTA.
//   kMH = 1: cluster tile 256 x 256, 6 stages, TMEM double-buffered (epilogue fully overlapped).
//   kMH = 2: cluster tile 512 x 256, 4 stages, all 512 TMEM columns hold one tile.  Per flop it
//            moves 25 % fewer operand bytes into shared memory (48 KB per 256 x 256 x 64 per CTA
//            instead of 32 KB per 128 x 256 x 64); shared-memory bandwidth (TMA fill + UMMA operand
//            reads, ~128 B/clk/SM for the 1-half tile) is what bounds the tensor pipe at this size,
//            so this wins even though the epilogue (8 warps) is no longer hidden behind the next
//            tile's MMAs -- only behind its TMA prefetch.
// kMode: 0 = TN  (A K-major [M,K], B K-major [N,K]),            bf16 output
//        1 = wgrad (A MN-major [rows,Mw], B MN-major [rows,Nw]),  fp32 += output
//        2 = NN  (A K-major [M,K], B MN-major [K,N]),             bf16 output -- dgrad dX = dY . W
//            straight from the row-major weight, no transposed copy
template <int kMode, int kMH, bool kDynamic>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64 + 128 * kMH, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, const Gemm2Params p) {
  constexpr bool kWgrad = kMode == 1;
  constexpr bool kAmn = kMode == 1, kBmn = kMode != 0;
  constexpr int kS = (kMH == 1) ? 6 : 4;
  constexpr uint32_t kABytes = uint32_t(kMH) * k2HalfBytes;
  constexpr uint32_t kStageBytes = kABytes + k2HalfBytes;
  constexpr int kAccBufs = (kMH == 1) ? 2 : 1;
  constexpr int kRowsPerCta = 128 * kMH;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  uint8_t* c_stage = tiles + kS * kStageBytes;                         // 2 x 16 KB (bf16 output path)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(c_stage + (kWgrad ? 0 : 2 * k2StoreBytes));
  uint64_t* empty_bar = full_bar + kS;
  uint64_t* tmem_full = empty_bar + kS;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  uint8_t* clc_resp = reinterpret_cast<uint8_t*>(tmem_ptr + 4);          // 16-byte aligned
  uint64_t* clc_full = reinterpret_cast<uint64_t*>(clc_resp + kClcSlots * 16);
  uint64_t* clc_empty = clc_full + kClcSlots;
  constexpr uint32_t kClcConsumers = 3 + 8 * kMH;   // 2 producers + MMA warp + epilogue warps of both CTAs

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  TileCursor cursor{clc_resp, clc_full, clc_empty, 0, 0u, num_clusters, num_tiles};

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (!kWgrad) tma_prefetch_desc(&tmap_c);
    for (int s = 0; s < kS; ++s) {
      mbar_init(&full_bar[s], 1);            // leader's expect_tx arrival; bytes come from both CTAs
      mbar_init(&empty_bar[s], 1);           // multicast tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);           // multicast tcgen05.commit
      mbar_init(&tmem_empty[a], 2 * 4 * kMH);   // one arrival per epilogue warp of either CTA
    }
    for (int i = 0; i < kClcSlots; ++i) {
      mbar_init(&clc_full[i], 1);            // armed remotely by the scheduler warp (expect_tx 16 B)
      mbar_init(&clc_empty[i], kClcConsumers);
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc_pair(tmem_ptr);
  tc_fence_before();
  cluster_sync_all();                        // barriers of BOTH CTAs are initialised before any remote use
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cluster_id; t >= 0; t = cursor.next(kDynamic, t, true)) {
        int m_blk, n_blk;
        tile_coords(p, t, m_blk, n_blk);
        const int a_row = (m_blk * 2 + int(rank)) * kRowsPerCta;   // this CTA's rows of the M extent
        const int b_row = n_blk * k2Tile + int(rank) * 128;        // this CTA's half of the N extent
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = tiles + stage * kStageBytes;
          uint8_t* b_dst = a_dst + kABytes;
          const uint32_t bar = map_to_cta(smem_u32(&full_bar[stage]), 0);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);
          // K-major operand: one [128 rows x 64 k] box per 128-row block.  MN-major operand: boxes of
          // [64 reduction rows x 64 columns]; a 128-wide extent is two boxes 8 KB apart.
          if (!kAmn) {
#pragma unroll
            for (int h = 0; h < kMH; ++h)
              tma_load_2d_pair(a_dst + h * k2HalfBytes, &tmap_a, bar, kb * 64, a_row + h * 128);
          } else {
#pragma unroll
            for (int c = 0; c < 2 * kMH; ++c)
              tma_load_2d_pair(a_dst + c * 8192, &tmap_a, bar, a_row + c * 64, kb * 64);
          }
          if (!kBmn) {
            tma_load_2d_pair(b_dst, &tmap_b, bar, kb * 64, b_row);
          } else {
            tma_load_2d_pair(b_dst, &tmap_b, bar, b_row, kb * 64);
            tma_load_2d_pair(b_dst + 8192, &tmap_b, bar, b_row + 64, kb * 64);
          }
          if (++stage == kS) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    // The whole warp runs the loop (warp-uniform control flow keeps the descriptor arithmetic on the
    // uniform datapath); one elected lane issues.  The issue loop must stay well under the 128-cycle
    // cost of one 256 x 256 x 16 UMMA or the tensor pipe starves on instruction issue.
    if (leader) {
      const uint32_t idesc = umma_idesc_bf16(256, 256, kAmn, kBmn);
      // K-major tile: 128-byte rows along K, advance 32 B per 16-wide k step.  MN-major tile: one
      // 128-byte row per k index, 64-wide MN chunks LBO = 8 KB apart, advance 2048 B per k step.
      const uint64_t desc_a0 = kAmn ? umma_smem_desc(smem_u32(tiles), 8192, 1024)
                                    : umma_smem_desc(smem_u32(tiles), 16, 1024);
      const uint64_t desc_b0 = kBmn ? umma_smem_desc(smem_u32(tiles + kABytes), 8192, 1024)
                                    : umma_smem_desc(smem_u32(tiles + kABytes), 16, 1024);
      constexpr uint64_t kStageStep = kStageBytes >> 4;        // descriptor address field is in 16-byte units
      constexpr uint64_t kHalfStep = k2HalfBytes >> 4;
      constexpr uint64_t kAStep = kAmn ? (2048 >> 4) : (32 >> 4);
      constexpr uint64_t kBStep = kBmn ? (2048 >> 4) : (32 >> 4);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = cluster_id; t >= 0; t = cursor.next(kDynamic, t, lane == 0)) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + uint32_t(acc) * 256;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t a_desc = desc_a0 + uint64_t(stage) * kStageStep;
            const uint64_t b_desc = desc_b0 + uint64_t(stage) * kStageStep;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
#pragma unroll
              for (int h = 0; h < kMH; ++h)
                umma_bf16_pair(d_tmem + uint32_t(h) * 256, a_desc + uint64_t(h) * kHalfStep + uint64_t(k) * kAStep,
                               b_desc + uint64_t(k) * kBStep, idesc, (k == 0) ? uint32_t(kb != 0) : 1u);
            }
            umma_commit_pair(&empty_bar[stage]);
            if (kb == p.num_k_blocks - 1) umma_commit_pair(&tmem_full[acc]);
          }
          __syncwarp();
          if (++stage == kS) { stage = 0; phase ^= 1; }
        }
        if (++acc == kAccBufs) { acc = 0; acc_phase ^= 1; }
      }
    } else if (kDynamic) {
      // ===================== tile scheduler (the otherwise idle warp 1 of CTA 1) =====================
      int slot = 0;
      uint32_t phase = 0;
      while (true) {
        mbar_wait(&clc_empty[slot], phase ^ 1);              // every consumer has read the old response
        if (lane < 2) mbar_expect_tx_cluster(map_to_cta(smem_u32(&clc_full[slot]), uint32_t(lane)), 16);
        __syncwarp();
        if (elect_one()) clc_try_cancel(clc_resp + slot * 16, &clc_full[slot]);
        mbar_wait(&clc_full[slot], phase);
        const int nt = clc_decode(clc_resp + slot * 16);
        if (++slot == kClcSlots) { slot = 0; phase ^= 1; }
        if (nt < 0) break;                                   // nothing left: the consumers see the same answer
      }
    }
  } else {
    // ===================== epilogue warps (both CTAs, own rows) =====================
    // kMH groups of 4 warps; group h drains accumulator half h (rows h*128 .. h*128+127 of the CTA)
    const int quad = warp & 3;
    const int h = (warp - 2) >> 2;
    const int r = quad * 32 + lane;                    // row inside the 128-row half
    const bool storer = (((warp - 2) & 3) == 0 && lane == 0);
    const int bar_id = 1 + h;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = cluster_id; t >= 0; t = cursor.next(kDynamic, t, lane == 0)) {
      int m_blk, n_blk;
      tile_coords(p, t, m_blk, n_blk);
      const int row0 = (m_blk * 2 + int(rank)) * kRowsPerCta + h * 128;
      const int n0 = n_blk * k2Tile;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (uint32_t(quad * 32) << 16) + uint32_t(kMH == 1 ? acc : h) * 256;
      if (!kWgrad) {
        // Phase A: drain the accumulator into registers (bias, ReLU, bf16 pack: 256 columns = 128
        // registers) and hand the TMEM back to the MMA warp at once; phase B (staging + TMA stores)
        // then overlaps the next tile's MMAs even when TMEM holds a single tile (kMH == 2).
        uint32_t packed[128];
#pragma unroll
        for (int c8 = 0; c8 < 8; ++c8) {
          uint32_t raw[32];
          tmem_ld_32x32(t_row + uint32_t(c8 * 32), raw);
          tmem_ld_wait();
          float v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
          if (p.bias != nullptr) {
            const float4* b4 = reinterpret_cast<const float4*>(p.bias + n0 + c8 * 32);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 bb = __ldg(b4 + q);
              v[q * 4] += bb.x; v[q * 4 + 1] += bb.y; v[q * 4 + 2] += bb.z; v[q * 4 + 3] += bb.w;
            }
          }
          if (p.relu) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) packed[c8 * 16 + i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&tmem_empty[acc]), 0));
        // Phase B
#pragma unroll
        for (int j = 0; j < 4; ++j) {                  // 64-column blocks of the 256-wide tile
          // kMH == 1: two staging buffers alternate; kMH == 2: each group owns one buffer
          uint8_t* cbuf = c_stage + (kMH == 1 ? (j & 1) : h) * k2StoreBytes;
          if (storer) {
            if (kMH == 1) tma_store_wait_read_le1();
            else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          }
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
          uint8_t* crow = cbuf + uint32_t(r) * 128;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const uint32_t piece = uint32_t(q) ^ uint32_t(r & 7);
            *reinterpret_cast<uint4*>(crow + (piece << 4)) =
                make_uint4(packed[j * 32 + q * 4], packed[j * 32 + q * 4 + 1], packed[j * 32 + q * 4 + 2],
                           packed[j * 32 + q * 4 + 3]);
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
          if (storer) {
            tma_store_2d_b(&tmap_c, cbuf, n0 + j * 64, row0);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      } else {
        const int m = row0 + r;
#pragma unroll 1
        for (int c = 0; c < 256; c += 32) {
          uint32_t raw[32];
          tmem_ld_32x32(t_row + uint32_t(c), raw);
          tmem_ld_wait();
          if (m < p.M) {
            float* dst = p.out_f32 + size_t(m) * p.ldo + n0 + c;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              if (n0 + c + q * 4 < p.N)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + q * 4),
                             "f"(__uint_as_float(raw[q * 4])), "f"(__uint_as_float(raw[q * 4 + 1])),
                             "f"(__uint_as_float(raw[q * 4 + 2])), "f"(__uint_as_float(raw[q * 4 + 3]))
                             : "memory");
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&tmem_empty[acc]), 0));
      }
      if (++acc == kAccBufs) { acc = 0; acc_phase ^= 1; }
    }
    if (!kWgrad && storer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  cluster_sync_all();                         // neither CTA may exit (or free TMEM) while the pair is active
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base);
  }
}

// Tile schedule of the pair kernels.  Static (74 persistent clusters, tile += 74) wastes the least
// at the tail when the GEMM owns the GPU; dynamic (cluster launch control) is for GEMMs that share
// the SMs with an overlapped all-reduce.  rdp_gemm_set_dynamic_tiles() / RAYDP_B200_DYNAMIC_TILES=1.
static int g_dynamic_tiles = -1;
static bool dynamic_tiles() {
  if (g_dynamic_tiles < 0) {
    const char* e = getenv("RAYDP_B200_DYNAMIC_TILES");
    g_dynamic_tiles = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  return g_dynamic_tiles == 1;
}
void gemm2_set_dynamic_tiles(int on) { g_dynamic_tiles = on ? 1 : 0; }

template <int kMode, int kMH>
static int launch_pair(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, Gemm2Params p,
                       cudaStream_t stream) {
  RDP_SMEM_ATTR_ONCE_PER_DEVICE((gemm_pair_kernel<kMode, kMH, true>), 227 * 1024);
  RDP_SMEM_ATTR_ONCE_PER_DEVICE((gemm_pair_kernel<kMode, kMH, false>), 227 * 1024);
  constexpr bool kWgrad = kMode == 1;
  constexpr int kS = (kMH == 1) ? 6 : 4;
  const size_t smem = size_t(kS) * (kMH + 1) * k2HalfBytes + (kWgrad ? 0 : 2 * k2StoreBytes) + 512 + 1024;
  p.num_m_tiles = (p.M + 256 * kMH - 1) / (256 * kMH);
  p.num_n_tiles = (p.N + k2Tile - 1) / k2Tile;
  // ~74 tiles in flight: a group of g row-blocks shares g A-panels and about 74/g B-panels
  int g = 8 / kMH;
  if (p.num_n_tiles < 10) g = (74 + p.num_n_tiles - 1) / p.num_n_tiles;
  if (g > p.num_m_tiles) g = p.num_m_tiles;
  p.group_m = g < 1 ? 1 : g;
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  if (dynamic_tiles()) {
    // one cluster per tile; resident clusters steal the not-yet-launched ones (cluster launch control)
    gemm_pair_kernel<kMode, kMH, true><<<2 * tiles, 64 + 128 * kMH, smem, stream>>>(ta, tb, tc, p);
  } else {
    int clusters = num_sms() / 2;
    if (tiles < clusters) clusters = tiles;
    gemm_pair_kernel<kMode, kMH, false><<<2 * clusters, 64 + 128 * kMH, smem, stream>>>(ta, tb, tc, p);
  }
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// 512-row cluster tiles once there are enough of them to fill the machine; 256-row tiles otherwise
static int pick_mh(int mh, int M, int N) {
  if (mh == 1 || mh == 2) return mh;
  const long tiles512 = long((M + 511) / 512) * ((N + 255) / 256);
  return tiles512 >= 2L * (num_sms() / 2) ? 2 : 1;
}

int gemm2_tn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K,
             const float* bias, int relu, int mh, cudaStream_t stream) {
  RDP_REQUIRE(N % k2Tile == 0, "gemm2_tn: N=%d must be a multiple of 256", N);
  RDP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "gemm2_tn: pitches must be multiples of 8");
  Gemm2Params p;
  p.M = M; p.N = N; p.K = K;
  p.num_k_blocks = (K + 63) / 64;
  p.bias = bias; p.relu = relu; p.out_f32 = nullptr; p.ldo = 0;
  CUtensorMap ta, tb, tc;
  int rc = make_tmap_2d_bf16(&ta, a, M, K, lda, 128, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tb, b, N, K, ldb, 128, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tc, out, M, N, ldc, 128, 64);
  if (rc) return rc;
  if (pick_mh(mh, M, N) == 2) return launch_pair<0, 2>(ta, tb, tc, p, stream);
  return launch_pair<0, 1>(ta, tb, tc, p, stream);
}

// out[M, N] (bf16) = a[M, K] . b[K, N]   (b row-major: the MN-major B operand)
int gemm2_nn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K, int mh,
             cudaStream_t stream) {
  RDP_REQUIRE(N % k2Tile == 0, "gemm2_nn: N=%d must be a multiple of 256", N);
  RDP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "gemm2_nn: pitches must be multiples of 8");
  Gemm2Params p;
  p.M = M; p.N = N; p.K = K;
  p.num_k_blocks = (K + 63) / 64;
  p.bias = nullptr; p.relu = 0; p.out_f32 = nullptr; p.ldo = 0;
  CUtensorMap ta, tb, tc;
  int rc = make_tmap_2d_bf16(&ta, a, M, K, lda, 128, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tb, b, K, N, ldb, 64, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tc, out, M, N, ldc, 128, 64);
  if (rc) return rc;
  if (pick_mh(mh, M, N) == 2) return launch_pair<2, 2>(ta, tb, tc, p, stream);
  return launch_pair<2, 1>(ta, tb, tc, p, stream);
}

int gemm2_wgrad(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw, int Nw,
                int mh, cudaStream_t stream) {
  RDP_REQUIRE(ldg % 8 == 0 && ldx % 8 == 0 && ldo % 4 == 0 && Nw % 4 == 0,
              "gemm2_wgrad: pitches must be multiples of 8 (inputs) / 4 (output)");
  Gemm2Params p;
  p.M = Mw; p.N = Nw; p.K = rows;
  p.num_k_blocks = (rows + 63) / 64;
  p.bias = nullptr; p.relu = 0; p.out_f32 = out; p.ldo = ldo;
  CUtensorMap tg, tx;
  int rc = make_tmap_2d_bf16(&tg, g, rows, Mw, ldg, 64, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tx, x, rows, Nw, ldx, 64, 64);
  if (rc) return rc;
  // default: 256-row tiles.  The fp32 red.add epilogue is only free while it hides behind the next
  // tile's MMAs, i.e. with the double-buffered accumulator (measured: 1343 vs 1198 TFLOP/s).
  if (mh == 2) return launch_pair<1, 2>(tg, tx, tx, p, stream);
  return launch_pair<1, 1>(tg, tx, tx, p, stream);
}

}  // namespace rdp

extern "C" {

// 1: pair kernels use hardware work stealing (cluster launch control) instead of the static persistent
// tile loop -- for GEMMs that run concurrently with an overlapped all-reduce kernel.  Process-wide.
int rdp_gemm_set_dynamic_tiles(int on) {
  rdp::gemm2_set_dynamic_tiles(on);
  return 0;
}

// Explicit entry points of the CTA-pair kernels (rdp_gemm_tn / rdp_gemm_wgrad pick them
// automatically for large problems; these let tests and benchmarks force the path).
// mh: 0 = choose, 1 = 256-row cluster tiles (double-buffered TMEM), 2 = 512-row cluster tiles.
int rdp_gemm2_tn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K,
                 const float* bias, int relu, int mh, void* stream) {
  return rdp::gemm2_tn(a, lda, b, ldb, out, ldc, M, N, K, bias, relu, mh, static_cast<cudaStream_t>(stream));
}

int rdp_gemm2_nn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K, int mh,
                 void* stream) {
  return rdp::gemm2_nn(a, lda, b, ldb, out, ldc, M, N, K, mh, static_cast<cudaStream_t>(stream));
}

int rdp_gemm2_wgrad(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw, int Nw,
                    int mh, void* stream) {
  return rdp::gemm2_wgrad(g, ldg, x, ldx, out, ldo, rows, Mw, Nw, mh, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
