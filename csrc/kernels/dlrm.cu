// K4 / K9 and the DLRM loss (workload B of SURVEY.md section 2.6: reference
// examples/pytorch_dlrm.ipynb cell 9 -- MultiTableEmbeddings, DotInteraction -- and cell 14).
//
// K4  embedding lookup fused with the model-parallel all-to-all.  Tables are sharded table-wise
//     across ranks.  The owner of a table reads the *requesting* rank's index column straight out
//     of that rank's memory, gathers the rows from its local shard and stores them directly into
//     the requesting rank's feature slab [B, 1+T, D] -- index dispatch, gather and combine are one
//     kernel and there are no staging buffers (`idx` / `feats` pointers of remote ranks are
//     NVLink peer mappings; with one rank they are simply local).  Backward mirrors it: the owner
//     pulls the gradient rows from every rank's dfeats slab and applies the sparse SGD update
//     with fp32 atomics (duplicate indices accumulate), so the embedding gradient never exists as
//     a tensor (reference: nn.Embedding(sparse=True) + DDP, which NCCL cannot all-reduce at all).
// K9  dot interaction: per sample the 27x27 Gram of the feature vectors, lower triangle (351
//     values, row-major i > j like torch.tril_indices(offset=-1)), concatenated after the dense
//     vector and zero-padded to 384 -- written in the layout the top-MLP GEMM reads through TMA.
#include <stdlib.h>
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

constexpr int kMaxSrc = 8;
constexpr int kMaxLocalTables = 32;

struct EmbParams {
  float* table[kMaxLocalTables];        // local shard: table[t] is [rows_t, D] fp32
  int slot[kMaxLocalTables];            // global feature slot (1-based row inside the slab)
  int col[kMaxLocalTables];             // column of this table inside the index matrix
  int n_tables;
  const int* idx[kMaxSrc];              // idx[s]: rank s' index matrix [B_s, idx_ld] int32
  __nv_bfloat16* feats[kMaxSrc];        // feats[s]: rank s' slab [B_s, S, D] (fwd: written, bwd: read)
  int batch[kMaxSrc];
  int n_src;
  int idx_ld, S, D;
  float scale;                          // bwd: -lr * grad_scale
  const float* lr_ptr;                  // bwd, optional: lr read from the device (CUDA-graph friendly);
  float grad_scale;                     //   then scale = -(*lr_ptr) * grad_scale
};

// one warp per (source rank, sample, local table) row; lanes cover D in steps of 32
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const __grid_constant__ EmbParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  int64_t total = 0;
  for (int s = 0; s < p.n_src; ++s) total += int64_t(p.batch[s]) * p.n_tables;
  for (int64_t w = warp; w < total; w += nwarps) {
    int64_t rem = w;
    int s = 0;
    while (rem >= int64_t(p.batch[s]) * p.n_tables) { rem -= int64_t(p.batch[s]) * p.n_tables; ++s; }
    const int b = int(rem / p.n_tables), t = int(rem % p.n_tables);
    const int row = p.idx[s][size_t(b) * p.idx_ld + p.col[t]];           // peer load (tiny)
    const float* src = p.table[t] + size_t(row) * p.D;
    __nv_bfloat16* dst = p.feats[s] + (size_t(b) * p.S + p.slot[t]) * p.D; // peer store
    for (int d = lane; d < p.D; d += 32) dst[d] = __float2bfloat16_rn(__ldg(src + d));
  }
}

__global__ void __launch_bounds__(256)
embedding_bwd_sgd_kernel(const __grid_constant__ EmbParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  int64_t total = 0;
  for (int s = 0; s < p.n_src; ++s) total += int64_t(p.batch[s]) * p.n_tables;
  const float scale = p.lr_ptr != nullptr ? -(*p.lr_ptr) * p.grad_scale : p.scale;
  for (int64_t w = warp; w < total; w += nwarps) {
    int64_t rem = w;
    int s = 0;
    while (rem >= int64_t(p.batch[s]) * p.n_tables) { rem -= int64_t(p.batch[s]) * p.n_tables; ++s; }
    const int b = int(rem / p.n_tables), t = int(rem % p.n_tables);
    const int row = p.idx[s][size_t(b) * p.idx_ld + p.col[t]];
    const __nv_bfloat16* g = p.feats[s] + (size_t(b) * p.S + p.slot[t]) * p.D;   // peer load
    float* dst = p.table[t] + size_t(row) * p.D;
    for (int d = lane; d < p.D; d += 32) atomicAdd(dst + d, scale * __bfloat162float(g[d]));
  }
}

// Vector variants for D % 4 == 0 with D / 4 a divisor of 32 (D = 32 -> 8 lanes per row, 4 rows per
// warp): every lane moves 4 elements (16-byte table access, 8-byte slab access) and a warp keeps 4
// independent idx -> row -> slab chains in flight instead of one, which is what these latency-bound
// gathers need (one row per warp took 37.7 us for 213 k rows; the row chains were serialised).
__device__ __forceinline__ bool emb_locate(const EmbParams& p, int64_t w, int& s, int& b, int& t) {
  int64_t rem = w;
  s = 0;
  while (s < p.n_src && rem >= int64_t(p.batch[s]) * p.n_tables) { rem -= int64_t(p.batch[s]) * p.n_tables; ++s; }
  if (s >= p.n_src) return false;
  b = int(rem / p.n_tables);
  t = int(rem % p.n_tables);
  return true;
}

__global__ void __launch_bounds__(256)
embedding_fwd_vec_kernel(const __grid_constant__ EmbParams p, int lanes_per_row, int64_t total) {
  const int lane = threadIdx.x & 31;
  const int rows_per_warp = 32 / lanes_per_row;
  const int sub = lane / lanes_per_row, part = lane % lanes_per_row;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t w0 = warp * rows_per_warp; w0 < total; w0 += nwarps * rows_per_warp) {
    const int64_t w = w0 + sub;
    int s, b, t;
    if (w >= total || !emb_locate(p, w, s, b, t)) continue;
    const int row = p.idx[s][size_t(b) * p.idx_ld + p.col[t]];
    const float4 v = __ldg(reinterpret_cast<const float4*>(p.table[t] + size_t(row) * p.D) + part);
    uint2* dst = reinterpret_cast<uint2*>(p.feats[s] + (size_t(b) * p.S + p.slot[t]) * p.D) + part;
    *dst = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
}

__global__ void __launch_bounds__(256)
embedding_bwd_sgd_vec_kernel(const __grid_constant__ EmbParams p, int lanes_per_row, int64_t total) {
  const int lane = threadIdx.x & 31;
  const int rows_per_warp = 32 / lanes_per_row;
  const int sub = lane / lanes_per_row, part = lane % lanes_per_row;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  const float scale = p.lr_ptr != nullptr ? -(*p.lr_ptr) * p.grad_scale : p.scale;
  for (int64_t w0 = warp * rows_per_warp; w0 < total; w0 += nwarps * rows_per_warp) {
    const int64_t w = w0 + sub;
    int s, b, t;
    if (w >= total || !emb_locate(p, w, s, b, t)) continue;
    const int row = p.idx[s][size_t(b) * p.idx_ld + p.col[t]];
    const uint2 g = *(reinterpret_cast<const uint2*>(p.feats[s] + (size_t(b) * p.S + p.slot[t]) * p.D) + part);
    const float2 g0 = unpack_bf16x2(g.x), g1 = unpack_bf16x2(g.y);
    float* dst = p.table[t] + size_t(row) * p.D + part * 4;
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(scale * g0.x), "f"(scale * g0.y),
                 "f"(scale * g1.x), "f"(scale * g1.y) : "memory");
  }
}

// ---- dot interaction ---------------------------------------------------------------------------
constexpr int kMaxS = 32;
// tensor-core path for D = 32 (dot_interaction_mma.cu); the kernels below serve every other shape
bool dot_interaction_mma_ok(int S, int D, int ld_out, const void* a, const void* b, const void* c);
int dot_interaction_fwd_mma(const void* feats, void* out, int B, int S, int ld_out, cudaStream_t stream);
int dot_interaction_bwd_mma(const void* feats, const void* dout, void* dfeats, int B, int S, int ld_out,
                            cudaStream_t stream);

__global__ void __launch_bounds__(128)
dot_interaction_fwd_kernel(const __nv_bfloat16* __restrict__ feats, __nv_bfloat16* __restrict__ out, int B,
                           int S, int D, int ld_out) {
  extern __shared__ float sh[];            // [S][D + 1]
  const int n_pairs = S * (S - 1) / 2;
  const int width = D + n_pairs;
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    const __nv_bfloat16* f = feats + size_t(b) * S * D;
    for (int i = threadIdx.x; i < S * D; i += blockDim.x) sh[(i / D) * (D + 1) + (i % D)] = __bfloat162float(f[i]);
    __syncthreads();
    __nv_bfloat16* o = out + size_t(b) * ld_out;
    for (int c = threadIdx.x; c < ld_out; c += blockDim.x) {
      float v = 0.f;
      if (c < D) {
        v = sh[c];                                       // dense vector = feature row 0
      } else if (c < width) {
        const int q = c - D;                             // pair (i, j), i > j, q = i (i - 1) / 2 + j
        int i = int((1.f + sqrtf(1.f + 8.f * float(q))) * 0.5f);
        while (i * (i - 1) / 2 > q) --i;
        while ((i + 1) * i / 2 <= q) ++i;
        const int j = q - i * (i - 1) / 2;
        const float* fi = sh + i * (D + 1);
        const float* fj = sh + j * (D + 1);
#pragma unroll 8
        for (int d = 0; d < D; ++d) v = fmaf(fi[d], fj[d], v);
      }
      o[c] = __float2bfloat16_rn(v);
    }
    __syncthreads();
  }
}

// dfeats[i, :] = sum_j dpair[i, j] * feats[j, :]  (+ the dense passthrough on row 0).
// 8 warps x 4 feature rows each cover S <= 32 rows in one pass; lanes run over D.  A thread keeps its 4
// row accumulators in registers, reads feats[j, d] once per j and the 4 dpair rows as warp-uniform
// 128-bit broadcasts (dpair rows are padded to SP = 32 floats), i.e. 0.5 shared-memory loads per FMA
// instead of the 2 of the element-per-thread formulation this replaced (which was LDS-bound).
constexpr int kDotSP = 32;                 // padded pair-matrix row length
// (4 blocks per SM: the per-sample chain load -> sync -> compute -> store is latency bound, so several
//  samples must be in flight per SM; a fully unrolled j loop took 171 registers = 1 block per SM and was 2x slower)
__global__ void __launch_bounds__(256, 4)
dot_interaction_bwd_kernel(const __nv_bfloat16* __restrict__ feats, const __nv_bfloat16* __restrict__ dout,
                           __nv_bfloat16* __restrict__ dfeats, int B, int S, int D, int ld_out) {
  extern __shared__ __align__(16) float sh[];   // dpair [32][32] (symmetric, zero diagonal/padding) | feats [32][D + 1]
  float* sp = sh;
  float* sf = sh + kDotSP * kDotSP;
  const int n_pairs = S * (S - 1) / 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kDotSP * (D + 1); i += blockDim.x) sf[i] = 0.f;     // rows >= S stay zero
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    const __nv_bfloat16* f = feats + size_t(b) * S * D;
    const __nv_bfloat16* go = dout + size_t(b) * ld_out;
    __syncthreads();                                   // previous sample fully consumed
    for (int i = threadIdx.x; i < S * D; i += blockDim.x) sf[(i / D) * (D + 1) + (i % D)] = __bfloat162float(f[i]);
    for (int i = threadIdx.x; i < kDotSP * kDotSP; i += blockDim.x) sp[i] = 0.f;
    __syncthreads();
    for (int q = threadIdx.x; q < n_pairs; q += blockDim.x) {
      int i = int((1.f + sqrtf(1.f + 8.f * float(q))) * 0.5f);
      while (i * (i - 1) / 2 > q) --i;
      while ((i + 1) * i / 2 <= q) ++i;
      const int j = q - i * (i - 1) / 2;
      const float g = __bfloat162float(go[D + q]);
      sp[i * kDotSP + j] = g;
      sp[j * kDotSP + i] = g;
    }
    __syncthreads();
    __nv_bfloat16* df = dfeats + size_t(b) * S * D;
    const int i0 = warp * 4;                           // this warp's 4 feature rows
    if (i0 < S) {
      for (int d = lane; d < D; d += 32) {
        float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll 2
        for (int j4 = 0; j4 < kDotSP; j4 += 4) {
          const float f0 = sf[(j4 + 0) * (D + 1) + d], f1 = sf[(j4 + 1) * (D + 1) + d];
          const float f2 = sf[(j4 + 2) * (D + 1) + d], f3 = sf[(j4 + 3) * (D + 1) + d];
          const float4 p0 = *reinterpret_cast<const float4*>(sp + (i0 + 0) * kDotSP + j4);
          const float4 p1 = *reinterpret_cast<const float4*>(sp + (i0 + 1) * kDotSP + j4);
          const float4 p2 = *reinterpret_cast<const float4*>(sp + (i0 + 2) * kDotSP + j4);
          const float4 p3 = *reinterpret_cast<const float4*>(sp + (i0 + 3) * kDotSP + j4);
          v0 = fmaf(p0.x, f0, fmaf(p0.y, f1, fmaf(p0.z, f2, fmaf(p0.w, f3, v0))));
          v1 = fmaf(p1.x, f0, fmaf(p1.y, f1, fmaf(p1.z, f2, fmaf(p1.w, f3, v1))));
          v2 = fmaf(p2.x, f0, fmaf(p2.y, f1, fmaf(p2.z, f2, fmaf(p2.w, f3, v2))));
          v3 = fmaf(p3.x, f0, fmaf(p3.y, f1, fmaf(p3.z, f2, fmaf(p3.w, f3, v3))));
        }
        if (i0 == 0) v0 += __bfloat162float(go[d]);     // dense passthrough
        const float vs[4] = {v0, v1, v2, v3};
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (i0 + r < S) df[(i0 + r) * D + d] = __float2bfloat16_rn(vs[r]);
      }
    }
  }
}

// ---- BCE-with-logits on a strided logit column -------------------------------------------------
__global__ void __launch_bounds__(256)
bce_logits_kernel(const __nv_bfloat16* __restrict__ logits, int ld, const float* __restrict__ target,
                  __nv_bfloat16* __restrict__ dlogits, int ld_d, int width_d, float* accum, int B, float grad_scale,
                  float* dbias) {
  float s_loss = 0.f, s_corr = 0.f, s_cnt = 0.f, s_dl = 0.f;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < B; r += gridDim.x * blockDim.x) {
    const float y = __bfloat162float(logits[size_t(r) * ld]);
    const float t = target[r];
    s_loss += fmaxf(y, 0.f) - y * t + log1pf(__expf(-fabsf(y)));
    s_corr += ((y >= 0.f) == (t >= 0.5f)) ? 1.f : 0.f;
    s_cnt += 1.f;
    if (dlogits != nullptr) {
      const float sig = 1.f / (1.f + __expf(-y));
      __nv_bfloat16* d = dlogits + size_t(r) * ld_d;
      const __nv_bfloat16 dl = __float2bfloat16_rn((sig - t) * grad_scale);
      d[0] = dl;
      s_dl += __bfloat162float(dl);
      for (int c = 1; c < width_d; ++c) d[c] = __float2bfloat16_rn(0.f);
    }
  }
  s_loss = warp_sum(s_loss); s_corr = warp_sum(s_corr); s_cnt = warp_sum(s_cnt); s_dl = warp_sum(s_dl);
  __shared__ float red[8][4];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { red[wid][0] = s_loss; red[wid][1] = s_corr; red[wid][2] = s_cnt; red[wid][3] = s_dl; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, c = 0.f, n = 0.f, dl = 0.f;
    for (int w = 0; w < 8; ++w) { a += red[w][0]; c += red[w][1]; n += red[w][2]; dl += red[w][3]; }
    if (dbias != nullptr) atomicAdd(dbias, dl);
    atomicAdd(&accum[0], a);
    atomicAdd(&accum[3], n);
    atomicAdd(&accum[4], c);
    atomicAdd(&accum[5], a / float(B));
  }
}

static int fill_emb(EmbParams& p, float* const* tables, const int* slots, const int* cols, int n_tables,
                    const int* const* idx, void* const* feats, const int* batch, int n_src, int idx_ld, int S, int D) {
  RDP_REQUIRE(n_tables >= 0 && n_tables <= kMaxLocalTables, "embedding: at most %d local tables (got %d)",
              kMaxLocalTables, n_tables);
  RDP_REQUIRE(n_src >= 1 && n_src <= kMaxSrc, "embedding: n_src=%d (max %d)", n_src, kMaxSrc);
  for (int t = 0; t < n_tables; ++t) { p.table[t] = tables[t]; p.slot[t] = slots[t]; p.col[t] = cols[t]; }
  p.n_tables = n_tables;
  for (int s = 0; s < n_src; ++s) {
    p.idx[s] = idx[s];
    p.feats[s] = static_cast<__nv_bfloat16*>(feats[s]);
    p.batch[s] = batch[s];
  }
  p.n_src = n_src; p.idx_ld = idx_ld; p.S = S; p.D = D; p.scale = 0.f; p.lr_ptr = nullptr; p.grad_scale = 1.f;
  return 0;
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_embedding_fwd(float* const* tables, const int* slots, const int* cols, int n_tables, const int* const* idx,
                      void* const* feats, const int* batch, int n_src, int idx_ld, int S, int D, void* stream) {
  if (n_tables == 0) return 0;
  EmbParams p;
  int rc = fill_emb(p, tables, slots, cols, n_tables, idx, feats, batch, n_src, idx_ld, S, D);
  if (rc) return rc;
  int64_t rows = 0;
  for (int s = 0; s < n_src; ++s) rows += int64_t(batch[s]) * n_tables;
  const int lpr = D / 4;
  if (D % 4 == 0 && lpr >= 1 && lpr <= 32 && 32 % lpr == 0) {
    const int64_t per_block = 8 * (32 / lpr);
    int grid = int((rows + per_block - 1) / per_block);
    if (grid > num_sms() * 16) grid = num_sms() * 16;
    embedding_fwd_vec_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p, lpr, rows);
    RDP_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  int grid = int((rows + 7) / 8);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  if (grid < 1) grid = 1;
  embedding_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_embedding_bwd_sgd(float* const* tables, const int* slots, const int* cols, int n_tables,
                          const int* const* idx, void* const* dfeats, const int* batch, int n_src, int idx_ld,
                          int S, int D, float lr, float grad_scale, const float* lr_ptr, void* stream) {
  if (n_tables == 0) return 0;
  EmbParams p;
  int rc = fill_emb(p, tables, slots, cols, n_tables, idx, dfeats, batch, n_src, idx_ld, S, D);
  if (rc) return rc;
  p.scale = -lr * grad_scale;
  p.lr_ptr = lr_ptr;
  p.grad_scale = grad_scale;
  int64_t rows = 0;
  for (int s = 0; s < n_src; ++s) rows += int64_t(batch[s]) * n_tables;
  const int lpr = D / 4;
  if (D % 4 == 0 && lpr >= 1 && lpr <= 32 && 32 % lpr == 0) {
    const int64_t per_block = 8 * (32 / lpr);
    int grid = int((rows + per_block - 1) / per_block);
    if (grid > num_sms() * 16) grid = num_sms() * 16;
    embedding_bwd_sgd_vec_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p, lpr, rows);
    RDP_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  int grid = int((rows + 7) / 8);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  if (grid < 1) grid = 1;
  embedding_bwd_sgd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_dot_interaction_fwd(const void* feats, void* out, int B, int S, int D, int ld_out, void* stream) {
  RDP_REQUIRE(S <= kMaxS && D <= 128 && ld_out >= D + S * (S - 1) / 2, "dot_interaction: bad shape S=%d D=%d ld=%d", S, D, ld_out);
  if (dot_interaction_mma_ok(S, D, ld_out, feats, out, out) && !getenv("RAYDP_B200_DOT_SCALAR"))
    return dot_interaction_fwd_mma(feats, out, B, S, ld_out, static_cast<cudaStream_t>(stream));
  int grid = B < num_sms() * 8 ? B : num_sms() * 8;
  dot_interaction_fwd_kernel<<<grid, 128, S * (D + 1) * sizeof(float), static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(feats), static_cast<__nv_bfloat16*>(out), B, S, D, ld_out);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_dot_interaction_bwd(const void* feats, const void* dout, void* dfeats, int B, int S, int D, int ld_out,
                            void* stream) {
  RDP_REQUIRE(S <= kMaxS && D <= 128, "dot_interaction: bad shape S=%d D=%d", S, D);
  if (dot_interaction_mma_ok(S, D, ld_out, feats, dout, dfeats) && !getenv("RAYDP_B200_DOT_SCALAR"))
    return dot_interaction_bwd_mma(feats, dout, dfeats, B, S, ld_out, static_cast<cudaStream_t>(stream));
  int grid = B < num_sms() * 4 ? B : num_sms() * 4;      // 4 resident blocks per SM (64 registers x 256 threads)
  const size_t smem = (kDotSP * kDotSP + kDotSP * (D + 1)) * sizeof(float);
  dot_interaction_bwd_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(feats), static_cast<const __nv_bfloat16*>(dout),
      static_cast<__nv_bfloat16*>(dfeats), B, S, D, ld_out);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_bce_logits(const void* logits, int ld, const float* target, void* dlogits, int ld_d, int width_d,
                   float* accum, int B, float grad_scale, float* dbias, void* stream) {
  int grid = (B + 255) / 256;
  if (grid > num_sms() * 2) grid = num_sms() * 2;
  bce_logits_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(logits), ld, target, static_cast<__nv_bfloat16*>(dlogits), ld_d, width_d,
      accum, B, grad_scale, dbias);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// One flag round between all ranks on `stream` (used between the phases of the embedding exchange).
// Implemented in csrc/comm/p2p_allreduce.cu.

}  // extern "C"
