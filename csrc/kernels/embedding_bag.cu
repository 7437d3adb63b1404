// K4 v2: pooled embedding lookups fused with the model-parallel all-to-all, and a DETERMINISTIC sparse
// update with row-wise optimizers (reference: examples/pytorch_dlrm.ipynb
// cell 9 keeps every table replicated under DDP -- nn.Embedding(sparse=True), bag size 1, plain SGD).
//
// What this adds over csrc/kernels/dlrm.cu (bag size 1, raw fp32 atomics, SGD):
//   * bags: table t contributes `len[t]` index columns per sample (multi-hot, fixed length, negative =
//     padding); the owner pools (sum or mean) while it gathers, so one D-vector per (sample, table) crosses
//     NVLink whatever the bag size;
//   * index dispatch by VECTOR loads: every rank publishes its indices TRANSPOSED ([column, sample]), so the
//     owner of a table reads the samples' indices of one requesting rank as 16-byte peer loads (4 samples
//     each) instead of one 4-byte peer load per row;
//   * uneven batches: the number of valid samples of every requesting rank is read from that rank's memory
//     (`batch_ptr`), so the last, shorter batch of an epoch needs no host round trip and fits a CUDA graph;
//   * duplicates: the rows a step touches are first given a slot in a compact accumulator (claim pass over the
//     indices only), then gradient rows are COMBINED per table row in 64-bit fixed point (integer adds are
//     associative, so the sum does not depend on the order the atomics land in: runs are bit-reproducible),
//     then every touched row is updated exactly once -- which row-wise Adagrad and Adam need to be correct
//     at all -- by SGD, row-wise Adagrad (one accumulator per row) or lazy Adam.  The accumulator holds one
//     row per index of a step (all local tables share it), not the tables.
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

constexpr int kBagMaxSrc = 8;
constexpr int kBagMaxTables = 32;
constexpr double kFixScale = 4294967296.0;          // 2^32: resolution 2.3e-10, range +-2.1e9

struct BagParams {
  float* table[kBagMaxTables];            // local shard: [rows_t, D] fp32
  int slot[kBagMaxTables];                // feature row inside the slab (1-based; 0 is the dense vector)
  int col[kBagMaxTables];                 // first index column of the table
  int len[kBagMaxTables];                 // bag length (index columns of the table)
  int n_tables;
  const int* idx[kBagMaxSrc];             // idx[s]: rank s' TRANSPOSED indices [n_cols, ldb] int32
  __nv_bfloat16* feats[kBagMaxSrc];       // fwd: slab to write; bwd: gradient slab to read; [cap, S, D]
  int batch[kBagMaxSrc];                  // capacity (upper bound) of rank s' batch
  const int* batch_ptr[kBagMaxSrc];       // optional: actual batch of rank s, in rank s' memory
  int n_src, ldb, S, D;
  int mean;                               // 0 sum, 1 mean over the valid entries of the bag
  // backward / update
  unsigned long long* acc;                // [slots, D] fixed-point gradient accumulators (zero between steps)
  int* flag[kBagMaxTables];               // [rows_t] 0 = untouched, else 1 + accumulator slot of the row
  int2* list;                             // [slots] touched rows of all local tables: list[slot] = (table, row)
  int* count;                             // [1] number of touched rows (zeroed by the host wrapper)
  float grad_scale;
};

__device__ __forceinline__ int valid_batch(const BagParams& p, int s) {
  int b = p.batch[s];
  if (p.batch_ptr[s] != nullptr) {
    const int actual = *reinterpret_cast<const volatile int*>(p.batch_ptr[s]);
    if (actual < b) b = actual;
  }
  return b < 0 ? 0 : b;
}

// work item = (source rank s, local table t, group of 4 consecutive samples); one sub-warp of D/4 lanes each
__device__ __forceinline__ bool bag_locate(const BagParams& p, int64_t w, int& s, int& t, int& b0) {
  int64_t rem = w;
  for (s = 0; s < p.n_src; ++s) {
    const int64_t groups = (int64_t(p.batch[s]) + 3) / 4 * p.n_tables;
    if (rem < groups) break;
    rem -= groups;
  }
  if (s >= p.n_src) return false;
  const int64_t per_table = (int64_t(p.batch[s]) + 3) / 4;
  t = int(rem / per_table);
  b0 = int(rem % per_table) * 4;
  return true;
}

__global__ void __launch_bounds__(256)
embedding_bag_fwd_kernel(const __grid_constant__ BagParams p, int lanes_per_row, int64_t total) {
  const int lane = threadIdx.x & 31;
  const int items_per_warp = 32 / lanes_per_row;
  const int sub = lane / lanes_per_row, part = lane % lanes_per_row;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t w0 = warp * items_per_warp; w0 < total; w0 += nwarps * items_per_warp) {
    const int64_t w = w0 + sub;
    int s, t, b0;
    if (w >= total || !bag_locate(p, w, s, t, b0)) continue;
    const int nb = valid_batch(p, s);
    if (b0 >= nb) continue;
    float4 sum[4];
    int cnt[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 4; ++k) sum[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int l = 0; l < p.len[t]; ++l) {
      // one 16-byte peer load: the indices of 4 samples for this (table, bag position)
      const int4 rows = *reinterpret_cast<const int4*>(p.idx[s] + size_t(p.col[t] + l) * p.ldb + b0);
      const int r[4] = {rows.x, rows.y, rows.z, rows.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (b0 + k < nb && r[k] >= 0) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(p.table[t] + size_t(r[k]) * p.D) + part);
          sum[k].x += v.x; sum[k].y += v.y; sum[k].z += v.z; sum[k].w += v.w;
          ++cnt[k];
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (b0 + k >= nb) break;
      const float f = (p.mean && cnt[k] > 1) ? 1.f / float(cnt[k]) : 1.f;
      uint2* dst = reinterpret_cast<uint2*>(p.feats[s] + (size_t(b0 + k) * p.S + p.slot[t]) * p.D) + part;
      *dst = make_uint2(pack_bf16x2(sum[k].x * f, sum[k].y * f), pack_bf16x2(sum[k].z * f, sum[k].w * f));
    }
  }
}

// Phase 0: claim.  One thread per (source, table, 4 samples): the first thread to see a row gives it the next
// accumulator slot.  Only indices are read.  Slots come from ONE counter, bumped once per warp (the lanes' claim
// counts are prefix-summed with shuffles): per-row atomics on a shared counter would serialise in L2.
__global__ void __launch_bounds__(256)
embedding_bag_claim_kernel(const __grid_constant__ BagParams p, int64_t total) {
  const int lane = threadIdx.x & 31;
  const int64_t first = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) - lane;
  for (int64_t w0 = first; w0 < total; w0 += int64_t(gridDim.x) * blockDim.x) {     // warp-uniform trip count
    const int64_t w = w0 + lane;
    int s = 0, t = 0, b0 = 0, nb = 0;
    bool live = w < total && bag_locate(p, w, s, t, b0);
    if (live) { nb = valid_batch(p, s); live = b0 < nb; }
    // pass 1: try to claim; remember the wins as a bit per (bag position, sample), 16 positions at a time
    const int len = live ? p.len[t] : 0;
    int max_len = len;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) max_len = max(max_len, __shfl_xor_sync(0xffffffffu, max_len, o));
    for (int l0 = 0; l0 < max_len; l0 += 16) {
      unsigned long long wins = 0ull;
      int nwin = 0;
      for (int l = l0; l < min(len, l0 + 16); ++l) {
        const int4 rows = *reinterpret_cast<const int4*>(p.idx[s] + size_t(p.col[t] + l) * p.ldb + b0);
        const int r[4] = {rows.x, rows.y, rows.z, rows.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (b0 + k >= nb || r[k] < 0) continue;
          if (atomicCAS(p.flag[t] + r[k], 0, -1) == 0) { wins |= 1ull << ((l - l0) * 4 + k); ++nwin; }
        }
      }
      // exclusive prefix of nwin over the warp, one atomic for the warp's total
      int incl = nwin;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const int warp_total = __shfl_sync(0xffffffffu, incl, 31);
      int base = 0;
      if (lane == 31 && warp_total > 0) base = atomicAdd(p.count, warp_total);
      base = __shfl_sync(0xffffffffu, base, 31);
      int pos = base + incl - nwin;
      // pass 2: hand out the slots (the indices are re-read: L1/L2 hits)
      if (wins) {
        for (int l = l0; l < min(len, l0 + 16); ++l) {
          const unsigned q = unsigned(wins >> ((l - l0) * 4)) & 15u;
          if (!q) continue;
          const int4 rows = *reinterpret_cast<const int4*>(p.idx[s] + size_t(p.col[t] + l) * p.ldb + b0);
          const int r[4] = {rows.x, rows.y, rows.z, rows.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (q & (1u << k)) {
              p.list[pos] = make_int2(t, r[k]);
              p.flag[t][r[k]] = pos + 1;        // read by the combine kernel, i.e. after this kernel ended
              ++pos;
            }
          }
        }
      }
    }
  }
}

// Phase A: combine.  Pull every requesting rank's gradient rows and add them, in fixed point, into the
// accumulator slot of the table row they belong to.
__global__ void __launch_bounds__(256)
embedding_bag_bwd_combine_kernel(const __grid_constant__ BagParams p, int lanes_per_row, int64_t total) {
  const int lane = threadIdx.x & 31;
  const int items_per_warp = 32 / lanes_per_row;
  const int sub = lane / lanes_per_row, part = lane % lanes_per_row;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t w0 = warp * items_per_warp; w0 < total; w0 += nwarps * items_per_warp) {
    const int64_t w = w0 + sub;
    int s, t, b0;
    if (w >= total || !bag_locate(p, w, s, t, b0)) continue;
    const int nb = valid_batch(p, s);
    if (b0 >= nb) continue;
    float4 g[4];
    int cnt[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      g[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (b0 + k < nb) {
        const uint2 raw = *(reinterpret_cast<const uint2*>(p.feats[s] + (size_t(b0 + k) * p.S + p.slot[t]) * p.D) + part);
        const float2 lo = unpack_bf16x2(raw.x), hi = unpack_bf16x2(raw.y);
        g[k] = make_float4(lo.x * p.grad_scale, lo.y * p.grad_scale, hi.x * p.grad_scale, hi.y * p.grad_scale);
      }
    }
    if (p.mean) {
      for (int l = 0; l < p.len[t]; ++l) {
        const int4 rows = *reinterpret_cast<const int4*>(p.idx[s] + size_t(p.col[t] + l) * p.ldb + b0);
        cnt[0] += rows.x >= 0; cnt[1] += rows.y >= 0; cnt[2] += rows.z >= 0; cnt[3] += rows.w >= 0;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (cnt[k] > 1) { const float f = 1.f / float(cnt[k]); g[k].x *= f; g[k].y *= f; g[k].z *= f; g[k].w *= f; }
      }
    }
    // fixed point once per sample, not once per bag position
    long long q[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      q[k][0] = __double2ll_rn(double(g[k].x) * kFixScale); q[k][1] = __double2ll_rn(double(g[k].y) * kFixScale);
      q[k][2] = __double2ll_rn(double(g[k].z) * kFixScale); q[k][3] = __double2ll_rn(double(g[k].w) * kFixScale);
    }
    for (int l = 0; l < p.len[t]; ++l) {
      const int4 rows = *reinterpret_cast<const int4*>(p.idx[s] + size_t(p.col[t] + l) * p.ldb + b0);
      const int r[4] = {rows.x, rows.y, rows.z, rows.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (b0 + k >= nb || r[k] < 0) continue;
        const int slot = p.flag[t][r[k]] - 1;
        unsigned long long* dst = p.acc + size_t(slot) * p.D + part * 4;
#pragma unroll
        for (int j = 0; j < 4; ++j) atomicAdd(dst + j, static_cast<unsigned long long>(q[k][j]));
      }
    }
  }
}

struct BagUpdateParams {
  float* table[kBagMaxTables];
  int* flag[kBagMaxTables];
  float* state[kBagMaxTables];            // Adagrad: [rows_t]; Adam: m then v, [2, rows_t, D]; SGD: unused
  int64_t state_stride[kBagMaxTables];    // Adam: rows_t * D (offset of v)
  unsigned long long* acc;
  const int2* list;
  const int* count;
  int D;
  int kind;                               // 0 SGD, 1 row-wise Adagrad, 2 lazy Adam
  float lr, eps, beta1, beta2, weight_decay;
  const float* lr_ptr;                    // optional device scalars (CUDA-graph friendly)
  const float* step_ptr;
};

// Phase B: every touched row exactly once: read the combined gradient, clear the accumulator, update.
// One flat list over all local tables: the load is balanced whatever the tables' sizes are.
__global__ void __launch_bounds__(256)
embedding_bag_update_kernel(const __grid_constant__ BagUpdateParams p, int lanes_per_row) {
  const int lane = threadIdx.x & 31;
  const int rows_per_warp = 32 / lanes_per_row;
  const int sub = lane / lanes_per_row, part = lane % lanes_per_row;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  const float lr = p.lr_ptr != nullptr ? *p.lr_ptr : p.lr;
  const float step = p.step_ptr != nullptr ? *p.step_ptr : 1.f;
  const unsigned sub_mask = (lanes_per_row == 32) ? 0xffffffffu : (((1u << lanes_per_row) - 1u) << (sub * lanes_per_row));
  float bc1 = 1.f, bc2 = 1.f;
  if (p.kind == 2) { bc1 = 1.f - powf(p.beta1, step); bc2 = 1.f - powf(p.beta2, step); }
  const int n = *p.count;
  // Two rows per sub-warp and iteration, and every load of an iteration issued before its first store: the loop
  // is a chain of dependent DRAM round trips (list -> accumulator / table row), so memory-level parallelism is
  // what it runs on (ncu, before: 13 % of DRAM bandwidth, 336 cycles of scoreboard stall per issue).
  constexpr int kU = 2;
  for (int64_t i0 = warp * rows_per_warp * kU; i0 < n; i0 += nwarps * rows_per_warp * kU) {
    int64_t i[kU];
    bool live[kU];
    int2 tr[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      i[u] = i0 + u * rows_per_warp + sub;
      live[u] = i[u] < n;
      tr[u] = live[u] ? __ldg(p.list + i[u]) : make_int2(0, 0);
    }
    ulonglong2 a0[kU], a1[kU];
    float4 w4[kU];
    float* wptr[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const ulonglong2* a = reinterpret_cast<const ulonglong2*>(p.acc + size_t(i[u]) * p.D + part * 4);
      wptr[u] = p.table[tr[u].x] + size_t(tr[u].y) * p.D + part * 4;
      a0[u] = a1[u] = make_ulonglong2(0ull, 0ull);
      w4[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (live[u]) {
        a0[u] = a[0]; a1[u] = a[1];
        w4[u] = *reinterpret_cast<const float4*>(wptr[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int t = tr[u].x, row = tr[u].y;
      float g[4];
      g[0] = float(double(static_cast<long long>(a0[u].x)) * (1.0 / kFixScale));
      g[1] = float(double(static_cast<long long>(a0[u].y)) * (1.0 / kFixScale));
      g[2] = float(double(static_cast<long long>(a1[u].x)) * (1.0 / kFixScale));
      g[3] = float(double(static_cast<long long>(a1[u].y)) * (1.0 / kFixScale));
      if (live[u]) {
        ulonglong2* a = reinterpret_cast<ulonglong2*>(p.acc + size_t(i[u]) * p.D + part * 4);
        a[0] = make_ulonglong2(0ull, 0ull);
        a[1] = make_ulonglong2(0ull, 0ull);
        if (part == 0) p.flag[t][row] = 0;
      }
      float w[4] = {w4[u].x, w4[u].y, w4[u].z, w4[u].w};
      if (p.weight_decay != 0.f) {
#pragma unroll
        for (int k = 0; k < 4; ++k) g[k] += p.weight_decay * w[k];
      }
      if (p.kind == 1) {
        // row-wise Adagrad: ONE accumulator per row, fed with the mean squared gradient of the row
        float sq = g[0] * g[0] + g[1] * g[1] + g[2] * g[2] + g[3] * g[3];
        for (int o = lanes_per_row / 2; o > 0; o >>= 1) sq += __shfl_xor_sync(sub_mask, sq, o);
        float acc = 0.f;
        if (live[u]) {
          acc = p.state[t][row] + sq / float(p.D);
          if (part == 0) p.state[t][row] = acc;
        }
        const float inv = lr / (sqrtf(acc) + p.eps);
#pragma unroll
        for (int k = 0; k < 4; ++k) w[k] -= inv * g[k];
      } else if (p.kind == 2) {
        if (live[u]) {
          float4* mp = reinterpret_cast<float4*>(p.state[t] + size_t(row) * p.D + part * 4);
          float4* vp = reinterpret_cast<float4*>(p.state[t] + p.state_stride[t] + size_t(row) * p.D + part * 4);
          const float4 m4 = *mp, v4 = *vp;
          float m[4] = {m4.x, m4.y, m4.z, m4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            m[k] = p.beta1 * m[k] + (1.f - p.beta1) * g[k];
            v[k] = p.beta2 * v[k] + (1.f - p.beta2) * g[k] * g[k];
            w[k] -= (lr / bc1) * (m[k] / (sqrtf(v[k]) / sqrtf(bc2) + p.eps));
          }
          *mp = make_float4(m[0], m[1], m[2], m[3]);
          *vp = make_float4(v[0], v[1], v[2], v[3]);
        }
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) w[k] -= lr * g[k];
      }
      if (live[u]) *reinterpret_cast<float4*>(wptr[u]) = make_float4(w[0], w[1], w[2], w[3]);
    }
  }
}

static int fill_bag(BagParams& p, float* const* tables, const int* slots, const int* cols, const int* lens, int n_tables,
                    const int* const* idx, void* const* feats, const int* batch, const int* const* batch_ptr, int n_src,
                    int ldb, int S, int D, int mean) {
  RDP_REQUIRE(n_tables >= 0 && n_tables <= kBagMaxTables, "embedding_bag: at most %d local tables (got %d)",
              kBagMaxTables, n_tables);
  RDP_REQUIRE(n_src >= 1 && n_src <= kBagMaxSrc, "embedding_bag: 1..%d source ranks (got %d)", kBagMaxSrc, n_src);
  RDP_REQUIRE(D % 4 == 0 && D / 4 >= 1 && D / 4 <= 32 && 32 % (D / 4) == 0,
              "embedding_bag: D=%d must be 4, 8, 16, 32, 64 or 128", D);
  RDP_REQUIRE(ldb % 4 == 0, "embedding_bag: the pitch of the transposed index matrix (%d) must be a multiple of 4", ldb);
  memset(&p, 0, sizeof(p));
  for (int t = 0; t < n_tables; ++t) {
    p.table[t] = tables[t]; p.slot[t] = slots[t]; p.col[t] = cols[t]; p.len[t] = lens ? lens[t] : 1;
    RDP_REQUIRE(p.len[t] >= 1, "embedding_bag: bag length of table %d is %d", t, p.len[t]);
  }
  p.n_tables = n_tables;
  for (int s = 0; s < n_src; ++s) {
    p.idx[s] = idx[s]; p.feats[s] = static_cast<__nv_bfloat16*>(feats[s]); p.batch[s] = batch[s];
    p.batch_ptr[s] = batch_ptr ? batch_ptr[s] : nullptr;
    RDP_REQUIRE(batch[s] <= ldb, "embedding_bag: batch %d exceeds the index pitch %d", batch[s], ldb);
  }
  p.n_src = n_src; p.ldb = ldb; p.S = S; p.D = D; p.mean = mean;
  return 0;
}

static int64_t bag_items(const BagParams& p) {
  int64_t total = 0;
  for (int s = 0; s < p.n_src; ++s) total += (int64_t(p.batch[s]) + 3) / 4 * p.n_tables;
  return total;
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_embedding_bag_fwd(float* const* tables, const int* slots, const int* cols, const int* lens, int n_tables,
                          const int* const* idx_t, void* const* feats, const int* batch, const int* const* batch_ptr,
                          int n_src, int ldb, int S, int D, int mean, void* stream) {
  if (n_tables == 0) return 0;
  BagParams p;
  int rc = fill_bag(p, tables, slots, cols, lens, n_tables, idx_t, feats, batch, batch_ptr, n_src, ldb, S, D, mean);
  if (rc) return rc;
  const int lpr = D / 4;
  const int64_t total = bag_items(p);
  if (total == 0) return 0;
  const int64_t per_block = 8 * (32 / lpr);
  int grid = int((total + per_block - 1) / per_block);
  if (grid > num_sms() * 16) grid = num_sms() * 16;
  embedding_bag_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p, lpr, total);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_embedding_bag_bwd(float* const* tables, const int* slots, const int* cols, const int* lens, int n_tables,
                          const int* const* idx_t, void* const* dfeats, const int* batch, const int* const* batch_ptr,
                          int n_src, int ldb, int S, int D, int mean, float grad_scale,
                          unsigned long long* acc, long long acc_rows, int* const* flag, int* list, int* count,
                          float* const* state, const int64_t* state_stride, int kind, float lr, float eps, float beta1,
                          float beta2, float weight_decay, const float* lr_ptr, const float* step_ptr, void* stream) {
  if (n_tables == 0) return 0;
  RDP_REQUIRE(kind >= 0 && kind <= 2, "rdp_embedding_bag_bwd: optimizer kind %d (0 SGD, 1 row-wise Adagrad, 2 Adam)", kind);
  BagParams p;
  int rc = fill_bag(p, tables, slots, cols, lens, n_tables, idx_t, dfeats, batch, batch_ptr, n_src, ldb, S, D, mean);
  if (rc) return rc;
  long long most = 0;                      // indices a step can carry = upper bound of touched rows
  for (int s = 0; s < n_src; ++s)
    for (int t = 0; t < n_tables; ++t) most += (long long)batch[s] * p.len[t];
  RDP_REQUIRE(acc_rows >= most,
              "rdp_embedding_bag_bwd: the accumulator holds %lld rows, a step can touch %lld", acc_rows, most);
  BagUpdateParams u;
  memset(&u, 0, sizeof(u));
  for (int t = 0; t < n_tables; ++t) {
    p.flag[t] = flag[t];
    u.table[t] = tables[t]; u.flag[t] = flag[t];
    u.state[t] = state ? state[t] : nullptr;
    u.state_stride[t] = state_stride ? state_stride[t] : 0;
    RDP_REQUIRE(kind == 0 || u.state[t] != nullptr, "rdp_embedding_bag_bwd: optimizer state of table %d missing", t);
  }
  p.acc = acc; p.list = reinterpret_cast<int2*>(list); p.count = count; p.grad_scale = grad_scale;
  u.acc = acc; u.list = reinterpret_cast<const int2*>(list); u.count = count; u.D = D; u.kind = kind;
  u.lr = lr; u.eps = eps; u.beta1 = beta1; u.beta2 = beta2;
  u.weight_decay = weight_decay; u.lr_ptr = lr_ptr; u.step_ptr = step_ptr;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  RDP_CHECK_CUDA(cudaMemsetAsync(count, 0, sizeof(int), s));
  const int lpr = D / 4;
  const int64_t total = bag_items(p);
  if (total == 0) return 0;
  const int64_t per_block = 8 * (32 / lpr);
  int grid = int((total + per_block - 1) / per_block);
  if (grid > num_sms() * 16) grid = num_sms() * 16;
  int cgrid = int((total + 255) / 256);
  if (cgrid > num_sms() * 8) cgrid = num_sms() * 8;
  embedding_bag_claim_kernel<<<cgrid, 256, 0, s>>>(p, total);
  RDP_CHECK_CUDA(cudaGetLastError());
  embedding_bag_bwd_combine_kernel<<<grid, 256, 0, s>>>(p, lpr, total);
  RDP_CHECK_CUDA(cudaGetLastError());
  embedding_bag_update_kernel<<<num_sms() * 8, 256, 0, s>>>(u, lpr);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
