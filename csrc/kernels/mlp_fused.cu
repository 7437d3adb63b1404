// K7 / K8 and the glue around K6 for Linear -> ReLU -> BatchNorm1d chains (workload A of
// reference examples/pytorch_nyctaxi.py:69-92).
//
// Design: BatchNorm is never materialised.  The GEMM epilogue emits per-column sum / sum-sq of
// the ReLU output z_l (rdp_gemm_tn stats_mode 1); rdp_bn_fold turns them into (a, c) with
// y_l = a*z_l + c and folds that affine map into the *next* layer's weights
// (W' = W diag(a), b' = b + W c), so the next GEMM reads z_l directly.  Backward mirrors it:
// the dgrad GEMM epilogue emits sum(dy) and sum(dy*z) (stats_mode 2), rdp_bn_relu_bwd applies
// the BatchNorm + ReLU backward in one pass, and rdp_grad_fixup maps weight gradients taken
// against z back to gradients against y.  rdp_mlp_head fuses the last Linear (-> 1 output),
// the loss, its derivative and the metric sums (A5-A7: no per-step host sync).
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

// ------------------------------------------------------------------------------------------
// BN finalize + fold into the next Linear
// ------------------------------------------------------------------------------------------
struct BnFoldParams {
  const float* stats;   // [2, K]  sum(z), sum(z*z) over the batch  (null when use_running)
  float count;          // batch rows
  const float* gamma;   // [K]
  const float* beta;    // [K]
  float* running_mean;  // [K] updated in place when training
  float* running_var;   // [K]
  float momentum, eps;
  int use_running;      // eval mode
  float* a;             // [K] out: gamma * rstd
  float* c;             // [K] out: beta - mean * a
  float* mean;          // [K] out
  float* rstd;          // [K] out
  const float* w;       // [N, K] next layer master weight (fp32)
  const float* b;       // [N]
  int N, K;
  __nv_bfloat16* w_folded;    // [N, ldw]   W * diag(a)
  int ldw;
  __nv_bfloat16* w_folded_t;  // [K, ldwt]  transposed UNFOLDED weight for dgrad (may be null)
  int ldwt;
  float* b_folded;            // [N]
};

__global__ void bn_fold_kernel(const BnFoldParams p) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  // one block per output row n of the next layer (+ extra blocks are not needed: N >= 1)
  extern __shared__ float sh[];
  float* s_a = sh;
  float* s_c = sh + p.K;
  for (int k = threadIdx.x; k < p.K; k += blockDim.x) {
    float mean, var;
    if (p.use_running) {
      mean = p.running_mean[k];
      var = p.running_var[k];
    } else {
      mean = p.stats[k] / p.count;
      var = fmaxf(p.stats[p.K + k] / p.count - mean * mean, 0.f);
    }
    const float rstd = rsqrtf(var + p.eps);
    const float a = p.gamma[k] * rstd;
    const float c = p.beta[k] - mean * a;
    s_a[k] = a;
    s_c[k] = c;
    if (blockIdx.x == 0) {
      p.a[k] = a; p.c[k] = c; p.mean[k] = mean; p.rstd[k] = rstd;
      if (!p.use_running && p.running_mean != nullptr) {
        const float unbiased = p.count > 1.f ? var * p.count / (p.count - 1.f) : var;
        p.running_mean[k] = (1.f - p.momentum) * p.running_mean[k] + p.momentum * mean;
        p.running_var[k] = (1.f - p.momentum) * p.running_var[k] + p.momentum * unbiased;
      }
    }
  }
  __syncthreads();
  const int n = blockIdx.x;
  if (n >= p.N) return;
  float dot = 0.f;
  for (int k = threadIdx.x; k < p.K; k += blockDim.x) {
    const float w = p.w[size_t(n) * p.K + k];
    const __nv_bfloat16 wf = __float2bfloat16_rn(w * s_a[k]);
    p.w_folded[size_t(n) * p.ldw + k] = wf;
    // dgrad needs dL/dy (y = BN output), i.e. the *unfolded* weight, transposed to K-major
    if (p.w_folded_t != nullptr) p.w_folded_t[size_t(k) * p.ldwt + n] = __float2bfloat16_rn(w);
    dot += w * s_c[k];
  }
  __shared__ float red[32];
  dot = warp_sum(dot);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = dot;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x + 31) / 32 ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) p.b_folded[n] = p.b[n] + v;
  }
}

// fp32 master weight -> bf16 (optionally padded pitch) + optional transposed copy
__global__ void cast_weight_kernel(const float* __restrict__ w, int N, int K, __nv_bfloat16* out,
                                   int ldo, __nv_bfloat16* out_t, int ldt) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * K) return;
  const int n = idx / K, k = idx % K;
  const __nv_bfloat16 v = __float2bfloat16_rn(w[idx]);
  out[size_t(n) * ldo + k] = v;
  if (out_t != nullptr) out_t[size_t(k) * ldt + n] = v;
}

// fp32 [N, K] -> bf16 [N, ldo] and its transpose bf16 [K, ldt] through a 32 x 32 shared-memory tile:
// both stores are coalesced (the element-per-thread kernel above writes the transpose with a stride)
__global__ void __launch_bounds__(256) cast_weight_tiled_kernel(const float* __restrict__ w, int N, int K,
                                                                __nv_bfloat16* __restrict__ out, int ldo,
                                                                __nv_bfloat16* __restrict__ out_t, int ldt) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int n = n0 + ty + r * 8, k = k0 + tx;
    float v = 0.f;
    if (n < N && k < K) {
      v = w[size_t(n) * K + k];
      out[size_t(n) * ldo + k] = __float2bfloat16_rn(v);
    }
    tile[ty + r * 8][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + r * 8, n = n0 + tx;
    if (k < K && n < N) out_t[size_t(k) * ldt + n] = __float2bfloat16_rn(tile[tx][ty + r * 8]);
  }
}

// All layers of a model in ONE launch: weights as above plus the fp32 bias copied into its padded operand.
// A step of the hybrid DLRM refreshes 8 layers: 8 cast launches + 8 bias copies become one kernel.
constexpr int kCastMaxLayers = 16;
struct CastManyParams {
  const float* w[kCastMaxLayers];
  __nv_bfloat16* out[kCastMaxLayers];
  __nv_bfloat16* out_t[kCastMaxLayers];
  const float* bias[kCastMaxLayers];
  float* bias_out[kCastMaxLayers];
  int N[kCastMaxLayers], K[kCastMaxLayers], ldo[kCastMaxLayers], ldt[kCastMaxLayers];
  int first_block[kCastMaxLayers + 1];
  int n_layers;
};

__global__ void __launch_bounds__(256) cast_weight_many_kernel(const __grid_constant__ CastManyParams p) {
  __shared__ float tile[32][33];
  int l = 0;
  while (l + 1 < p.n_layers && int(blockIdx.x) >= p.first_block[l + 1]) ++l;
  const int N = p.N[l], K = p.K[l], ldo = p.ldo[l], ldt = p.ldt[l];
  const int kb = (K + 31) / 32;
  const int rel = blockIdx.x - p.first_block[l];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int k0 = (rel % kb) * 32, n0 = (rel / kb) * 32;
  const float* __restrict__ w = p.w[l];
  __nv_bfloat16* __restrict__ out = p.out[l];
  __nv_bfloat16* __restrict__ out_t = p.out_t[l];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int n = n0 + ty + r * 8, k = k0 + tx;
    float v = 0.f;
    if (n < N && k < K) {
      v = w[size_t(n) * K + k];
      out[size_t(n) * ldo + k] = __float2bfloat16_rn(v);
    }
    tile[ty + r * 8][tx] = v;
  }
  __syncthreads();
  if (out_t != nullptr) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int k = k0 + ty + r * 8, n = n0 + tx;
      if (k < K && n < N) out_t[size_t(k) * ldt + n] = __float2bfloat16_rn(tile[tx][ty + r * 8]);
    }
  }
  if (k0 == 0 && p.bias[l] != nullptr && threadIdx.x < 32 && n0 + int(threadIdx.x) < N)
    p.bias_out[l][n0 + threadIdx.x] = p.bias[l][n0 + threadIdx.x];
}

// contiguous fp32 -> bf16, 8 elements per thread (no transposed copy)
__global__ void cast_bf16_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int64_t n8) {
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n8; i += int64_t(gridDim.x) * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(w)[2 * i], b = reinterpret_cast<const float4*>(w)[2 * i + 1];
    reinterpret_cast<uint4*>(out)[i] = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w),
                                                  pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
}

// ------------------------------------------------------------------------------------------
// ReLU backward + bias gradient for generic Linear(+ReLU) layers: one pass over [M, N]
//   g_out = (act > 0) ? g : 0   (act = the layer's ReLU output; skipped when act == nullptr)
//   dbias[col] += sum over rows of g_out
// Block = 32 column lanes (8 bf16 columns each = 256 columns) x 8 row lanes.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) relu_bwd_bias_kernel(const __nv_bfloat16* __restrict__ g,
                                                            const __nv_bfloat16* __restrict__ act,
                                                            __nv_bfloat16* __restrict__ g_out, float* __restrict__ dbias,
                                                            int M, int N, int ld, int rows_per_block) {
  __shared__ float red[8][256];
  const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + cl) * 8;
  const int r_begin = blockIdx.y * rows_per_block;
  const int r_end = min(M, r_begin + rows_per_block);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col < N) {
    for (int r = r_begin + rl; r < r_end; r += 8) {
      const size_t off = size_t(r) * ld + col;
      uint4 gv = *reinterpret_cast<const uint4*>(g + off);
      uint32_t* gu = reinterpret_cast<uint32_t*>(&gv);
      if (act != nullptr) {
        const uint4 av = *reinterpret_cast<const uint4*>(act + off);
        const uint32_t* au = reinterpret_cast<const uint32_t*>(&av);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          // bf16 > 0  <=>  sign bit clear and magnitude non-zero
          const uint32_t lo = au[i] & 0xffffu, hi = au[i] >> 16;
          const uint32_t keep = (((lo & 0x8000u) == 0 && (lo & 0x7fffu) != 0) ? 0x0000ffffu : 0u) |
                                (((hi & 0x8000u) == 0 && (hi & 0x7fffu) != 0) ? 0xffff0000u : 0u);
          gu[i] &= keep;
        }
        *reinterpret_cast<uint4*>(g_out + off) = gv;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] += __uint_as_float(gu[i] << 16);
        acc[2 * i + 1] += __uint_as_float(gu[i] & 0xffff0000u);
      }
    }
  }
  if (dbias == nullptr) return;
#pragma unroll
  for (int i = 0; i < 8; ++i) red[rl][cl * 8 + i] = acc[i];
  __syncthreads();
  const int c = threadIdx.x;                 // 256 threads <-> 256 columns of the block
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) sum += red[j][c];
  const int gc = blockIdx.x * 256 + c;
  if (gc < N) atomicAdd(dbias + gc, sum);
}

// ------------------------------------------------------------------------------------------
// BatchNorm + ReLU backward (elementwise, with column sums of the result for the bias grad)
// ------------------------------------------------------------------------------------------
struct BnReluBwdParams {
  const __nv_bfloat16* dy;   // [B, N] grad w.r.t. the BN output
  const __nv_bfloat16* z;    // [B, N] ReLU output (BN input)
  __nv_bfloat16* dpre;       // [B, N] grad w.r.t. the Linear output (pre-ReLU)
  int B, N, ld;
  const float* stats2;       // [2, N] sum(dy), sum(dy*z)
  const float* mean;         // [N]
  const float* rstd;         // [N]
  const float* a;            // [N] gamma*rstd
  float* dgamma;             // [N] out (=)
  float* dbeta;              // [N] out (=)
  float* dbias;              // [N] out (+=, must be zeroed): column sums of dpre
  int has_bn;                // 0: plain ReLU backward (dpre = dy * (z > 0))
};

__global__ void __launch_bounds__(256)
bn_relu_bwd_kernel(const BnReluBwdParams p) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  // thread -> 8 consecutive columns (one 16-byte vector); blockDim.x = 256
  const int vec_per_row = p.N / 8;
  const int tcol = threadIdx.x % vec_per_row;
  const int trow = threadIdx.x / vec_per_row;
  const int rows_per_block = blockDim.x / vec_per_row;
  const int c0 = tcol * 8;
  float m1[8], m2[8], mu[8], rs[8], aa[8], acc[8];
  const float invB = 1.f / float(p.B);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    acc[j] = 0.f;
    if (p.has_bn) {
      const float sdy = p.stats2[c0 + j], sdyz = p.stats2[p.N + c0 + j];
      mu[j] = p.mean[c0 + j];
      rs[j] = p.rstd[c0 + j];
      aa[j] = p.a[c0 + j];
      const float sdyzh = rs[j] * (sdyz - mu[j] * sdy);     // sum(dy * zhat)
      m1[j] = sdy * invB;
      m2[j] = sdyzh * invB;
      if (blockIdx.x == 0 && trow == 0) {
        p.dgamma[c0 + j] = sdyzh;
        p.dbeta[c0 + j] = sdy;
      }
    }
  }
  // four rows per thread per trip: all eight 16-byte loads are in flight before the first use
  const int row_stride = gridDim.x * rows_per_block;
  for (int r0 = blockIdx.x * rows_per_block + trow; r0 < p.B; r0 += 4 * row_stride) {
    uint4 udy[4], uz[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + u * row_stride;
      if (r < p.B) {
        udy[u] = __ldg(reinterpret_cast<const uint4*>(p.dy + size_t(r) * p.ld + c0));
        uz[u] = __ldg(reinterpret_cast<const uint4*>(p.z + size_t(r) * p.ld + c0));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + u * row_stride;
      if (r >= p.B) break;
      const uint32_t dyw[4] = {udy[u].x, udy[u].y, udy[u].z, udy[u].w};
      const uint32_t zw[4] = {uz[u].x, uz[u].y, uz[u].z, uz[u].w};
      uint32_t ow[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 d = unpack_bf16x2(dyw[q]);
        const float2 zz = unpack_bf16x2(zw[q]);
        float o0, o1;
        if (p.has_bn) {
          const float zh0 = (zz.x - mu[2 * q]) * rs[2 * q];
          const float zh1 = (zz.y - mu[2 * q + 1]) * rs[2 * q + 1];
          o0 = zz.x > 0.f ? aa[2 * q] * (d.x - m1[2 * q] - zh0 * m2[2 * q]) : 0.f;
          o1 = zz.y > 0.f ? aa[2 * q + 1] * (d.y - m1[2 * q + 1] - zh1 * m2[2 * q + 1]) : 0.f;
        } else {
          o0 = zz.x > 0.f ? d.x : 0.f;
          o1 = zz.y > 0.f ? d.y : 0.f;
        }
        ow[q] = pack_bf16x2(o0, o1);
        const float2 rounded = unpack_bf16x2(ow[q]);
        acc[2 * q] += rounded.x;
        acc[2 * q + 1] += rounded.y;
      }
      *reinterpret_cast<uint4*>(p.dpre + size_t(r) * p.ld + c0) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
  }
  extern __shared__ float s_col[];   // [N]
  for (int i = threadIdx.x; i < p.N; i += blockDim.x) s_col[i] = 0.f;
  __syncthreads();
  // lanes that own the same columns (narrow layers: several rows per warp) combine first, so
  // the shared-memory atomics see at most one add per column per warp
  bool leader = true;
  if (vec_per_row < 32) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      if (o >= vec_per_row) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
      }
    }
    leader = (threadIdx.x & 31) < vec_per_row;
  }
  if (leader) {
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&s_col[c0 + j], acc[j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < p.N; i += blockDim.x)
    if (s_col[i] != 0.f) atomicAdd(&p.dbias[i], s_col[i]);
}

// ------------------------------------------------------------------------------------------
// head: y = (a*z + c) . w + b ; loss ; dloss ; metrics ; dy = dpred * w ; reductions
// ------------------------------------------------------------------------------------------
struct HeadParams {
  const __nv_bfloat16* z;   // [B, K] last hidden activation (ReLU output)
  int B, K, ld;
  const float* a;           // [K] BN fold of the last BatchNorm (null -> identity)
  const float* c;           // [K]
  const float* w;           // [K] last Linear weight (fp32 master)
  const float* bias;        // [1]
  const float* target;      // [B]
  int loss_kind;            // 0 MSE, 1 SmoothL1 (beta 1), 2 BCEWithLogits
  float grad_scale;         // 1/B for mean reduction (0 in eval mode: no backward outputs)
  float* pred;              // [B] out (may be null)
  float* accum;             // [8]: loss_sum, abs_err_sum, sq_err_sum, count, correct, -, -, -
  __nv_bfloat16* dy;        // [B, K] out: grad w.r.t. the BN output of the last hidden layer
  float* stats2;            // [2, K] out (+=): sum(dy), sum(dy*z)
  float* dw;                // [K] out (+=): grad of w
  float* db;                // [1] out (+=)
};

template <int kK>
__global__ void __launch_bounds__(256)
mlp_head_kernel(const HeadParams p) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  float wf[kK], cw = 0.f;          // folded weight w*a and constant c.w
#pragma unroll
  for (int k = 0; k < kK; ++k) {
    const float a = p.a ? p.a[k] : 1.f, c = p.c ? p.c[k] : 0.f;
    wf[k] = p.w[k] * a;
    cw += p.w[k] * c;
  }
  const float bias = p.bias[0] + cw;
  float s_loss = 0.f, s_abs = 0.f, s_sq = 0.f, s_cnt = 0.f, s_corr = 0.f, s_dp = 0.f;
  float s_dpz[kK];
#pragma unroll
  for (int k = 0; k < kK; ++k) s_dpz[k] = 0.f;
  const bool train = p.grad_scale != 0.f;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < p.B; r += gridDim.x * blockDim.x) {
    float zr[kK];
    const uint4* zp = reinterpret_cast<const uint4*>(p.z + size_t(r) * p.ld);
#pragma unroll
    for (int q = 0; q < kK / 8; ++q) {
      const uint4 u = __ldg(zp + q);
      const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
      zr[q * 8 + 0] = f0.x; zr[q * 8 + 1] = f0.y; zr[q * 8 + 2] = f1.x; zr[q * 8 + 3] = f1.y;
      zr[q * 8 + 4] = f2.x; zr[q * 8 + 5] = f2.y; zr[q * 8 + 6] = f3.x; zr[q * 8 + 7] = f3.y;
    }
    float y = bias;
#pragma unroll
    for (int k = 0; k < kK; ++k) y = fmaf(zr[k], wf[k], y);
    const float t = p.target[r];
    float loss, dl;
    if (p.loss_kind == 2) {                      // BCE with logits, numerically stable
      loss = fmaxf(y, 0.f) - y * t + log1pf(__expf(-fabsf(y)));
      const float sig = 1.f / (1.f + __expf(-y));
      dl = sig - t;
      s_corr += ((y >= 0.f) == (t >= 0.5f)) ? 1.f : 0.f;
    } else {
      const float e = y - t;
      if (p.loss_kind == 0) { loss = e * e; dl = 2.f * e; }
      else { const float ae = fabsf(e); loss = ae < 1.f ? 0.5f * e * e : ae - 0.5f; dl = ae < 1.f ? e : (e > 0.f ? 1.f : -1.f); }
    }
    const float e = y - t;
    s_loss += loss; s_abs += fabsf(e); s_sq += e * e; s_cnt += 1.f;
    if (p.pred) p.pred[r] = y;
    if (train) {
      const float dp = dl * p.grad_scale;
      s_dp += dp;
      uint4* dyp = reinterpret_cast<uint4*>(p.dy + size_t(r) * p.ld);
#pragma unroll
      for (int q = 0; q < kK / 8; ++q) {
        uint32_t o[4];
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const int k = q * 8 + h * 2;
          // dy is the gradient w.r.t. the (virtual) BN output: dpred * w (unfolded weight)
          o[h] = pack_bf16x2(dp * p.w[k], dp * p.w[k + 1]);
        }
        dyp[q] = make_uint4(o[0], o[1], o[2], o[3]);
      }
#pragma unroll
      for (int k = 0; k < kK; ++k) s_dpz[k] = fmaf(dp, zr[k], s_dpz[k]);
    }
  }
  // block reduction -> global atomics
  __shared__ float red[8][6 + kK];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float vals[6] = {s_loss, s_abs, s_sq, s_cnt, s_corr, s_dp};
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const float v = warp_sum(vals[i]);
    if (lane == 0) red[wid][i] = v;
  }
#pragma unroll
  for (int k = 0; k < kK; ++k) {
    const float v = warp_sum(s_dpz[k]);
    if (lane == 0) red[wid][6 + k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 6 + kK) {
    float v = 0.f;
    for (int w8 = 0; w8 < 8; ++w8) v += red[w8][threadIdx.x];
    const int i = threadIdx.x;
    if (i < 5) {
      atomicAdd(&p.accum[i], v);
      // accum[5] collects the batch-mean loss, so an epoch's "mean of per-batch means"
      // (reference torch/estimator.py:290) needs no per-step host read
      if (i == 0) atomicAdd(&p.accum[5], v / float(p.B));
    } else if (train) {
      if (i == 5) {
        atomicAdd(p.db, v);
        // sum over batch of dy[:, k] = w[k] * sum(dp); and dw picks up the c term
        for (int k = 0; k < kK; ++k) {
          atomicAdd(&p.stats2[k], p.w[k] * v);
          atomicAdd(&p.dw[k], (p.c ? p.c[k] : 0.f) * v);
        }
      } else {
        const int k = i - 6;
        atomicAdd(&p.stats2[kK + k], p.w[k] * v);            // sum(dy * z)
        atomicAdd(&p.dw[k], (p.a ? p.a[k] : 1.f) * v);       // dw = a * sum(dp * z) + c * sum(dp)
      }
    }
  }
}

// dW[n,k] = dWraw[n,k] * a[k] + db[n] * c[k]   (gradient taken against z mapped to y = a z + c)
__global__ void grad_fixup_kernel(float* dw, const float* db, const float* a, const float* c, int N, int K) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * K) return;
  const int n = idx / K, k = idx % K;
  dw[idx] = dw[idx] * a[k] + db[n] * c[k];
}

// ------------------------------------------------------------------------------------------
// fused optimizers over the flat parameter buffer
// ------------------------------------------------------------------------------------------
struct AdamParams {
  float* p; const float* g; float* m; float* v;
  int64_t n;
  float lr, beta1, beta2, eps, weight_decay;
  float bc1, bc2;        // 1 - beta1^t, 1 - beta2^t
  float grad_scale;      // e.g. 1/world_size
  int adamw;
  const float* lr_ptr;   // optional device scalars (CUDA-graph friendly): lr, step count
  const float* step_ptr;
  __nv_bfloat16* shadow; // optional bf16 copy of the updated parameters
};

__device__ __forceinline__ float adam_update(float p, float g, float& m, float& v, const AdamParams& a) {
  g *= a.grad_scale;
  if (a.weight_decay != 0.f) { if (a.adamw) p *= (1.f - a.lr * a.weight_decay); else g += a.weight_decay * p; }
  m = a.beta1 * m + (1.f - a.beta1) * g;
  v = a.beta2 * v + (1.f - a.beta2) * g * g;
  const float denom = sqrtf(v) / sqrtf(a.bc2) + a.eps;
  return p - (a.lr / a.bc1) * (m / denom);
}

// kVec: 4 elements per thread with 16-byte accesses (pointers 16-byte aligned; the host checks)
template <bool kVec>
__global__ void adam_kernel(AdamParams a) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  if (a.lr_ptr != nullptr) a.lr = *a.lr_ptr;
  if (a.step_ptr != nullptr) {
    const float t = *a.step_ptr;
    a.bc1 = 1.f - powf(a.beta1, t);
    a.bc2 = 1.f - powf(a.beta2, t);
  }
  const int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x, nth = int64_t(gridDim.x) * blockDim.x;
  int64_t tail_begin = 0;
  if (kVec) {
    const int64_t n4 = a.n / 4;
    for (int64_t i = tid; i < n4; i += nth) {
      float4 p4 = reinterpret_cast<float4*>(a.p)[i], m4 = reinterpret_cast<float4*>(a.m)[i];
      float4 v4 = reinterpret_cast<float4*>(a.v)[i];
      const float4 g4 = reinterpret_cast<const float4*>(a.g)[i];
      p4.x = adam_update(p4.x, g4.x, m4.x, v4.x, a);
      p4.y = adam_update(p4.y, g4.y, m4.y, v4.y, a);
      p4.z = adam_update(p4.z, g4.z, m4.z, v4.z, a);
      p4.w = adam_update(p4.w, g4.w, m4.w, v4.w, a);
      reinterpret_cast<float4*>(a.p)[i] = p4;
      reinterpret_cast<float4*>(a.m)[i] = m4;
      reinterpret_cast<float4*>(a.v)[i] = v4;
      if (a.shadow != nullptr)
        reinterpret_cast<uint2*>(a.shadow)[i] = make_uint2(pack_bf16x2(p4.x, p4.y), pack_bf16x2(p4.z, p4.w));
    }
    tail_begin = n4 * 4;
  }
  for (int64_t i = tail_begin + tid; i < a.n; i += nth) {
    float m = a.m[i], v = a.v[i];
    const float pn = adam_update(a.p[i], a.g[i], m, v, a);
    a.p[i] = pn;
    a.m[i] = m; a.v[i] = v;
    if (a.shadow != nullptr) a.shadow[i] = __float2bfloat16_rn(pn);
  }
}

struct SgdParams {
  float* p; const float* g; float* buf;   // momentum buffer or null
  int64_t n;
  float lr, momentum, dampening, weight_decay, grad_scale;
  int nesterov, first_step;
  const float* lr_ptr;
  const float* step_ptr;   // first_step = (*step_ptr <= 1)
  __nv_bfloat16* shadow;   // optional bf16 copy of the updated parameters
};

// BatchNorm backward WITHOUT a pass over the batch.  For a hidden layer l-1 (r = relu(pre), y = BN(r)) feeding
// Linear l, the two batch sums BN backward needs are linear functionals of quantities the step already has:
//   sum_b dy[b, j]        = sum_i db_l[i]  * W_l[i, j]              (db_l = column sums of dpre_l)
//   sum_b dy[b, j] r[b,j] = sum_i W_l[i,j] * G_l[i, j]              (G_l = dpre_l^T r: the raw weight gradient)
// because dy = dpre_l . W_l.  So after wgrad_l this O(N K) kernel yields d(gamma), d(beta) of BN l-1 and the
// per-column constants of   dpre_{l-1} = (r > 0) ? alpha * dy + beta * r + delta : 0 ,  which the dgrad GEMM
// applies in its epilogue (rdp_gemm_tn_bnbwd): dy is never written, the separate elementwise
// BN+ReLU-backward pass over [B, width] (26 % of the step) is gone.  W is read from the bf16 transposed copy the
// dgrad GEMM multiplies by, so the sums match the dy that GEMM produces.  The same pass rewrites the raw
// gradient into the gradient of the UNFOLDED weight (y = a r + c was folded into W_l for the forward):
//   dW[i, j] = a[j] * G[i, j] + c[j] * db[i]         (what rdp_grad_fixup did as its own launch).
struct BnBwdCoefParams {
  float* dw; int ldw;                  // [N, K] raw G in, fixed gradient out
  const float* db;                     // [N]
  const __nv_bfloat16* w_t; int ldt;   // [K, N] bf16 transposed unfolded weight
  int N, K;
  const float* a; const float* c;      // [K] BN fold of layer l-1 (a = gamma * rstd)
  const float* mean; const float* rstd;
  float inv_count;
  float* dgamma; float* dbeta;         // [K] out (=)
  float* alpha; float* beta; float* delta;   // [K] out
};

constexpr int kCoefRows = 32;       // row phases per block: 32 columns x 32 phases = 1024 threads

__global__ void __launch_bounds__(32 * kCoefRows)
bn_bwd_coef_kernel(const BnBwdCoefParams p) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  __shared__ float red[2][kCoefRows][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + tx;
  float s1 = 0.f, t = 0.f;
  if (j < p.K) {
    const float aj = p.a[j], cj = p.c[j];
    for (int i = ty; i < p.N; i += kCoefRows) {
      const float g = p.dw[size_t(i) * p.ldw + j];
      const float w = __bfloat162float(p.w_t[size_t(j) * p.ldt + i]);
      const float dbi = p.db[i];
      s1 = fmaf(dbi, w, s1);
      t = fmaf(w, g, t);
      p.dw[size_t(i) * p.ldw + j] = fmaf(aj, g, cj * dbi);
    }
  }
  red[0][ty][tx] = s1;
  red[1][ty][tx] = t;
  __syncthreads();
  if (ty == 0 && j < p.K) {
    // fixed summation order: the result does not depend on scheduling
    float S1 = 0.f, T = 0.f;
#pragma unroll
    for (int k = 0; k < kCoefRows; ++k) { S1 += red[0][k][tx]; T += red[1][k][tx]; }
    const float mu = p.mean[j], rs = p.rstd[j], al = p.a[j];
    const float S2 = rs * (T - mu * S1);                  // sum_b dy * zhat
    p.dbeta[j] = S1;
    p.dgamma[j] = S2;
    const float be = -al * rs * S2 * p.inv_count;
    p.alpha[j] = al;
    p.beta[j] = be;
    p.delta[j] = -al * S1 * p.inv_count - be * mu;
  }
}

// The step's result (running loss / metric sums) goes to the host by a STORE from the device into
// page-locked, device-mapped host memory -- not by a cudaMemcpyAsync.  A device->host memcpy queued behind
// a training step sits in the copy queue until that step is done, and the host->device copies of the NEXT
// batches, issued after it, were observed to wait behind it (profiles/ingest_timeline.md): the per-step
// result read-back was what exposed the input copies.  A kernel store has no place in that queue.
__global__ void publish_host_kernel(const float* __restrict__ src, volatile float* dst_host, int n) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  // posted writes: no system-scope fence here -- waiting for the PCIe acknowledgement under a saturated
  // host->device stream cost up to 15 us per step; the host only reads after a stream synchronisation
  const int i = threadIdx.x;
  if (i < n) dst_host[i] = src[i];
}

__global__ void add_scalar_kernel(float* x, float v) { *x += v; }

__device__ __forceinline__ float sgd_update(float p, float g, float* buf, int64_t i, const SgdParams& s) {
  g *= s.grad_scale;
  if (s.weight_decay != 0.f) g += s.weight_decay * p;
  if (buf != nullptr && s.momentum != 0.f) {
    const float b = s.first_step ? g : s.momentum * buf[i] + (1.f - s.dampening) * g;
    buf[i] = b;
    g = s.nesterov ? g + s.momentum * b : b;
  }
  return p - s.lr * g;
}

template <bool kVec>
__global__ void sgd_kernel(SgdParams s) {
  pdl_trigger();      // see common.cuh: programmatic dependent launch
  pdl_wait();
  if (s.lr_ptr != nullptr) s.lr = *s.lr_ptr;
  if (s.step_ptr != nullptr) s.first_step = (*s.step_ptr <= 1.f) ? 1 : 0;
  const int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x, nth = int64_t(gridDim.x) * blockDim.x;
  int64_t tail_begin = 0;
  if (kVec) {
    const int64_t n4 = s.n / 4;
    for (int64_t i = tid; i < n4; i += nth) {
      float4 p4 = reinterpret_cast<float4*>(s.p)[i];
      const float4 g4 = reinterpret_cast<const float4*>(s.g)[i];
      p4.x = sgd_update(p4.x, g4.x, s.buf, 4 * i, s);
      p4.y = sgd_update(p4.y, g4.y, s.buf, 4 * i + 1, s);
      p4.z = sgd_update(p4.z, g4.z, s.buf, 4 * i + 2, s);
      p4.w = sgd_update(p4.w, g4.w, s.buf, 4 * i + 3, s);
      reinterpret_cast<float4*>(s.p)[i] = p4;
      if (s.shadow != nullptr)
        reinterpret_cast<uint2*>(s.shadow)[i] = make_uint2(pack_bf16x2(p4.x, p4.y), pack_bf16x2(p4.z, p4.w));
    }
    tail_begin = n4 * 4;
  }
  for (int64_t i = tail_begin + tid; i < s.n; i += nth) {
    const float pn = sgd_update(s.p[i], s.g[i], s.buf, i, s);
    s.p[i] = pn;
    if (s.shadow != nullptr) s.shadow[i] = __float2bfloat16_rn(pn);
  }
}

static inline bool aligned16(const void* a, const void* b, const void* c, const void* d, const void* e) {
  return ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
           reinterpret_cast<uintptr_t>(d)) & 15) == 0 && (reinterpret_cast<uintptr_t>(e) & 7) == 0;
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_bn_fold(const float* stats, float count, const float* gamma, const float* beta,
                float* running_mean, float* running_var, float momentum, float eps, int use_running,
                float* a, float* c, float* mean, float* rstd, const float* w, const float* b, int N, int K,
                void* w_folded, int ldw, void* w_folded_t, int ldwt, float* b_folded, void* stream) {
  RDP_REQUIRE(K > 0 && N > 0 && K <= 8192, "rdp_bn_fold: bad shape N=%d K=%d", N, K);
  BnFoldParams p{stats, count, gamma, beta, running_mean, running_var, momentum, eps, use_running,
                 a, c, mean, rstd, w, b, N, K, static_cast<__nv_bfloat16*>(w_folded), ldw,
                 static_cast<__nv_bfloat16*>(w_folded_t), ldwt, b_folded};
  const int threads = K >= 256 ? 256 : (K >= 128 ? 128 : 64);
  RDP_CHECK_CUDA(launch_pdl(bn_fold_kernel, dim3(N), dim3(threads), 2 * K * sizeof(float), static_cast<cudaStream_t>(stream), p));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_cast_weight(const float* w, int N, int K, void* out, int ldo, void* out_t, int ldt, void* stream) {
  const int total = N * K;
  if (out_t == nullptr && ldo == K && total % 8 == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0 &&
      (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
    const int64_t n8 = total / 8;
    int grid = int((n8 + 255) / 256);
    if (grid > num_sms() * 16) grid = num_sms() * 16;
    cast_bf16_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(w, static_cast<__nv_bfloat16*>(out), n8);
    RDP_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  if (out_t != nullptr) {
    dim3 grid((K + 31) / 32, (N + 31) / 32);
    cast_weight_tiled_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        w, N, K, static_cast<__nv_bfloat16*>(out), ldo, static_cast<__nv_bfloat16*>(out_t), ldt);
    RDP_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  cast_weight_kernel<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w, N, K, static_cast<__nv_bfloat16*>(out), ldo, static_cast<__nv_bfloat16*>(out_t), ldt);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_cast_weight_many(const float* const* w, const int* N, const int* K, void* const* out, const int* ldo,
                         void* const* out_t, const int* ldt, const float* const* bias, float* const* bias_out,
                         int n_layers, void* stream) {
  RDP_REQUIRE(n_layers >= 1 && n_layers <= kCastMaxLayers, "rdp_cast_weight_many: 1..%d layers (got %d)",
              kCastMaxLayers, n_layers);
  CastManyParams p;
  memset(&p, 0, sizeof(p));
  int blocks = 0;
  for (int l = 0; l < n_layers; ++l) {
    p.w[l] = w[l]; p.out[l] = static_cast<__nv_bfloat16*>(out[l]);
    p.out_t[l] = out_t ? static_cast<__nv_bfloat16*>(out_t[l]) : nullptr;
    p.bias[l] = bias ? bias[l] : nullptr; p.bias_out[l] = bias_out ? bias_out[l] : nullptr;
    p.N[l] = N[l]; p.K[l] = K[l]; p.ldo[l] = ldo[l]; p.ldt[l] = ldt ? ldt[l] : 0;
    p.first_block[l] = blocks;
    blocks += ((K[l] + 31) / 32) * ((N[l] + 31) / 32);
  }
  p.first_block[n_layers] = blocks;
  p.n_layers = n_layers;
  cast_weight_many_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_bn_relu_bwd(const void* dy, const void* z, void* dpre, int B, int N, int ld, const float* stats2,
                    const float* mean, const float* rstd, const float* a, float* dgamma, float* dbeta,
                    float* dbias, int has_bn, void* stream) {
  RDP_REQUIRE(N % 8 == 0 && N <= 2048 && ld % 8 == 0, "rdp_bn_relu_bwd: N=%d must be a multiple of 8 and <= 2048", N);
  RDP_REQUIRE(256 % (N / 8) == 0 || N / 8 > 256, "rdp_bn_relu_bwd: N/8=%d must divide 256", N / 8);
  BnReluBwdParams p{static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(z),
                    static_cast<__nv_bfloat16*>(dpre), B, N, ld, stats2, mean, rstd, a, dgamma, dbeta, dbias, has_bn};
  const int rows_per_block = 256 / (N / 8);
  int grid = (B + 4 * rows_per_block - 1) / (4 * rows_per_block);
  const int cap = num_sms() * 8;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  RDP_CHECK_CUDA(launch_pdl(bn_relu_bwd_kernel, dim3(grid), dim3(256), N * sizeof(float), static_cast<cudaStream_t>(stream), p));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_mlp_head(const void* z, int B, int K, int ld, const float* a, const float* c, const float* w,
                 const float* bias, const float* target, int loss_kind, float grad_scale, float* pred,
                 float* accum, void* dy, float* stats2, float* dw, float* db, void* stream) {
  RDP_REQUIRE(K == 16 || K == 32 || K == 64, "rdp_mlp_head: K=%d (supported: 16, 32, 64)", K);
  HeadParams p{static_cast<const __nv_bfloat16*>(z), B, K, ld, a, c, w, bias, target, loss_kind, grad_scale,
               pred, accum, static_cast<__nv_bfloat16*>(dy), stats2, dw, db};
  int grid = (B + 255) / 256;
  const int cap = num_sms() * 2;
  if (grid > cap) grid = cap;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (K == 16) RDP_CHECK_CUDA(launch_pdl(mlp_head_kernel<16>, dim3(grid), dim3(256), 0, s, p));
  else if (K == 32) RDP_CHECK_CUDA(launch_pdl(mlp_head_kernel<32>, dim3(grid), dim3(256), 0, s, p));
  else RDP_CHECK_CUDA(launch_pdl(mlp_head_kernel<64>, dim3(grid), dim3(256), 0, s, p));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_grad_fixup(float* dw, const float* db, const float* a, const float* c, int N, int K, void* stream) {
  const int total = N * K;
  RDP_CHECK_CUDA(launch_pdl(grad_fixup_kernel, dim3((total + 255) / 256), dim3(256), 0, static_cast<cudaStream_t>(stream), dw, db, a, c, N, K));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_bn_bwd_coef(float* dw, int ldw, const float* db, const void* w_t, int ldt, int N, int K, const float* a,
                    const float* c, const float* mean, const float* rstd, float count, float* dgamma,
                    float* dbeta, float* alpha, float* beta, float* delta, void* stream) {
  RDP_REQUIRE(N > 0 && K > 0 && count > 0.f, "rdp_bn_bwd_coef: bad shape N=%d K=%d", N, K);
  BnBwdCoefParams p{dw, ldw, db, static_cast<const __nv_bfloat16*>(w_t), ldt, N, K, a, c, mean, rstd,
                    1.f / count, dgamma, dbeta, alpha, beta, delta};
  RDP_CHECK_CUDA(launch_pdl(bn_bwd_coef_kernel, dim3((K + 31) / 32), dim3(32 * kCoefRows), 0, static_cast<cudaStream_t>(stream), p));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_publish_host(const float* src, float* host_pinned, int n, void* stream) {
  RDP_REQUIRE(n > 0 && n <= 32, "rdp_publish_host: n=%d (1..32 floats)", n);
  void* mapped = nullptr;
  RDP_CHECK_CUDA(cudaHostGetDevicePointer(&mapped, host_pinned, 0));
  RDP_CHECK_CUDA(launch_pdl(publish_host_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(stream), src, static_cast<volatile float*>(static_cast<float*>(mapped)), n));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_add_scalar(float* x, float v, void* stream) {
  add_scalar_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(x, v);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                  float eps, float weight_decay, int step, float grad_scale, int adamw,
                  const float* lr_ptr, const float* step_ptr, void* shadow, void* stream) {
  AdamParams a{p, g, m, v, n, lr, beta1, beta2, eps, weight_decay,
               1.f - powf(beta1, float(step)), 1.f - powf(beta2, float(step)), grad_scale, adamw,
               lr_ptr, step_ptr, static_cast<__nv_bfloat16*>(shadow)};
  const bool vec = aligned16(p, g, m, v, shadow) && n >= 4;
  int grid = int(((vec ? n / 4 : n) + 255) / 256);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  if (grid < 1) grid = 1;
  if (vec) RDP_CHECK_CUDA(launch_pdl(adam_kernel<true>, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), a));
  else RDP_CHECK_CUDA(launch_pdl(adam_kernel<false>, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), a));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_sgd_step(float* p, const float* g, float* buf, int64_t n, float lr, float momentum, float dampening,
                 float weight_decay, int nesterov, int first_step, float grad_scale,
                 const float* lr_ptr, const float* step_ptr, void* shadow, void* stream) {
  SgdParams s{p, g, buf, n, lr, momentum, dampening, weight_decay, grad_scale, nesterov, first_step,
              lr_ptr, step_ptr, static_cast<__nv_bfloat16*>(shadow)};
  const bool vec = aligned16(p, g, buf, nullptr, shadow) && n >= 4;
  int grid = int(((vec ? n / 4 : n) + 255) / 256);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  if (grid < 1) grid = 1;
  if (vec) RDP_CHECK_CUDA(launch_pdl(sgd_kernel<true>, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), s));
  else RDP_CHECK_CUDA(launch_pdl(sgd_kernel<false>, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), s));
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// g_out = relu'(act) * g (skipped when act is NULL), dbias[N] += column sums of the result.
int rdp_relu_bwd_bias(const void* g, const void* act, void* g_out, float* dbias, int M, int N, int ld, void* stream) {
  RDP_REQUIRE(N % 8 == 0 && ld % 8 == 0, "rdp_relu_bwd_bias: N=%d and ld=%d must be multiples of 8", N, ld);
  RDP_REQUIRE(act == nullptr || g_out != nullptr, "rdp_relu_bwd_bias: g_out missing");
  const int col_blocks = (N + 255) / 256;
  int row_blocks = (num_sms() * 8 + col_blocks - 1) / col_blocks;
  int rows_per_block = (M + row_blocks - 1) / row_blocks;
  rows_per_block = ((rows_per_block + 7) / 8) * 8;
  if (rows_per_block < 8) rows_per_block = 8;
  row_blocks = (M + rows_per_block - 1) / rows_per_block;
  dim3 grid(col_blocks, row_blocks);
  relu_bwd_bias_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(g), static_cast<const __nv_bfloat16*>(act),
      static_cast<__nv_bfloat16*>(g_out), dbias, M, N, ld, rows_per_block);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
