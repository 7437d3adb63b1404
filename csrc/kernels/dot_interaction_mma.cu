// K9 on the tensor cores: DLRM dot interaction for D = 32, S <= 32 (the reference's shape: 27 x 32,
// examples/pytorch_dlrm.ipynb cell 9 `DotInteraction`: torch.bmm + tril gather + cat + pad).
//
// One warp per sample.  The sample's feature matrix F [S, 32] (bf16, 1.7 KB) is staged once in shared memory;
//   forward : P = F F^T        (32 x 32 x 32, zero rows past S) -> packed strict lower triangle
//   backward: dF = G F          with G the symmetric matrix rebuilt from the packed pair gradients
// are 16 `mma.sync.m16n8k16` each, with fragments read by `ldmatrix` (the backward's B operand by
// `ldmatrix.trans`), instead of the 351 scalar 32-long dot products per sample of the element-per-thread
// formulation (measured 44 us forward / 68 us backward at B = 8192; the data is 20 MB).  The problem is a batch of
// 27 x 27 x 32 products: far below a tcgen05 tile (M = 128) and bound by the 20 MB it streams, so the warp-level
// MMA is the right instrument here; results leave through a shared-memory staging row as 16-byte stores.
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

constexpr int kDiD = 32;            // embedding width handled here
constexpr int kDiPitch = 40;        // bf16 elements per staged row (80 bytes: conflict-free ldmatrix phases)
constexpr int kDiWarps = 4;
constexpr int kDiMaxOut = 544;      // >= ld_out: 32 + 31 * 32 / 2 = 528, padded

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* smem_ptr) {
  const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(smem_ptr));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], const void* smem_ptr) {
  const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(smem_ptr));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
               "{%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct DiSmem {
  __nv_bfloat16 f[kDiWarps][32][kDiPitch];      // staged features (rows >= S stay zero)
  __nv_bfloat16 g[kDiWarps][32][kDiPitch];      // backward: symmetric pair-gradient matrix
  __nv_bfloat16 row[kDiWarps][kDiMaxOut];       // packed output row (fwd) / packed gradient row (bwd)
  unsigned char pi[512], pj[512];               // pair q -> (i, j), i > j
};

__device__ __forceinline__ void di_init(DiSmem& sm, int S) {
  for (int i = threadIdx.x; i < kDiWarps * 32 * kDiPitch; i += blockDim.x) {
    (&sm.f[0][0][0])[i] = __float2bfloat16(0.f);
    (&sm.g[0][0][0])[i] = __float2bfloat16(0.f);
  }
  for (int i = threadIdx.x; i < kDiWarps * kDiMaxOut; i += blockDim.x) (&sm.row[0][0])[i] = __float2bfloat16(0.f);
  for (int q = threadIdx.x; q < S * (S - 1) / 2; q += blockDim.x) {
    int i = int((1.f + sqrtf(1.f + 8.f * float(q))) * 0.5f);
    while (i * (i - 1) / 2 > q) --i;
    while ((i + 1) * i / 2 <= q) ++i;
    sm.pi[q] = (unsigned char)i;
    sm.pj[q] = (unsigned char)(q - i * (i - 1) / 2);
  }
  __syncthreads();
}

// copy one sample's [S, 32] bf16 rows (contiguous in global memory) into the padded staging tile
__device__ __forceinline__ void di_stage(__nv_bfloat16 (*dst)[kDiPitch], const __nv_bfloat16* src, int S, int lane) {
  for (int c = lane; c < S * 4; c += 32) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(src) + c);
    *reinterpret_cast<uint4*>(&dst[c >> 2][(c & 3) * 8]) = v;
  }
}

__global__ void __launch_bounds__(kDiWarps * 32)
dot_interaction_fwd_mma_kernel(const __nv_bfloat16* __restrict__ feats, __nv_bfloat16* __restrict__ out, int B, int S,
                               int ld_out) {
  __shared__ __align__(16) DiSmem sm;
  di_init(sm, S);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tq = lane & 3;
  __nv_bfloat16 (*sf)[kDiPitch] = sm.f[warp];
  __nv_bfloat16* so = sm.row[warp];
  for (int b = blockIdx.x * kDiWarps + warp; b < B; b += gridDim.x * kDiWarps) {
    di_stage(sf, feats + size_t(b) * S * kDiD, S, lane);
    __syncwarp();
    uint32_t a[2][2][4], bf[4][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) ldsm_x4(a[mt][ks], &sf[mt * 16 + (lane & 15)][ks * 16 + (lane >> 4) * 8]);
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) ldsm_x4(bf[nt], &sf[nt * 8 + (lane & 7)][(lane >> 3) * 8]);
    float c[2][4][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
        for (int r = 0; r < 4; ++r) c[mt][nt][r] = 0.f;
        mma_bf16_16816(c[mt][nt], a[mt][0], bf[nt][0], bf[nt][1]);
        mma_bf16_16816(c[mt][nt], a[mt][1], bf[nt][2], bf[nt][3]);
      }
    // packed row: [F row 0 | strict lower triangle, row-major | zero pad]
    so[lane] = sf[0][lane];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const int i = mt * 16 + g + (r >> 1) * 8, j = nt * 8 + tq * 2 + (r & 1);
          if (i < S && j < i) so[kDiD + i * (i - 1) / 2 + j] = __float2bfloat16_rn(c[mt][nt][r]);
        }
    __syncwarp();
    uint4* o = reinterpret_cast<uint4*>(out + size_t(b) * ld_out);
    for (int v = lane; v < ld_out / 8; v += 32) o[v] = *reinterpret_cast<const uint4*>(so + v * 8);
    __syncwarp();
  }
}

__global__ void __launch_bounds__(kDiWarps * 32)
dot_interaction_bwd_mma_kernel(const __nv_bfloat16* __restrict__ feats, const __nv_bfloat16* __restrict__ dout,
                               __nv_bfloat16* __restrict__ dfeats, int B, int S, int ld_out) {
  __shared__ __align__(16) DiSmem sm;
  di_init(sm, S);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tq = lane & 3;
  const int n_pairs = S * (S - 1) / 2;
  __nv_bfloat16 (*sf)[kDiPitch] = sm.f[warp];
  __nv_bfloat16 (*sg)[kDiPitch] = sm.g[warp];
  __nv_bfloat16* so = sm.row[warp];
  for (int b = blockIdx.x * kDiWarps + warp; b < B; b += gridDim.x * kDiWarps) {
    di_stage(sf, feats + size_t(b) * S * kDiD, S, lane);
    const uint4* go = reinterpret_cast<const uint4*>(dout + size_t(b) * ld_out);
    for (int v = lane; v < ld_out / 8; v += 32) *reinterpret_cast<uint4*>(so + v * 8) = __ldg(go + v);
    __syncwarp();
    for (int q = lane; q < n_pairs; q += 32) {       // G is symmetric with a zero diagonal
      const int i = sm.pi[q], j = sm.pj[q];
      const __nv_bfloat16 v = so[kDiD + q];
      sg[i][j] = v;
      sg[j][i] = v;
    }
    __syncwarp();
    uint32_t a[2][2][4], bf[4][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) ldsm_x4(a[mt][ks], &sg[mt * 16 + (lane & 15)][ks * 16 + (lane >> 4) * 8]);
    // B[k = j][n = d] = F[j][d]: rows are k, so the 8x8 blocks are transposed on load
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) ldsm_x4_trans(bf[nt], &sf[(lane >> 3) * 8 + (lane & 7)][nt * 8]);
    float c[2][4][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
        for (int r = 0; r < 4; ++r) c[mt][nt][r] = 0.f;
        mma_bf16_16816(c[mt][nt], a[mt][0], bf[nt][0], bf[nt][1]);
        mma_bf16_16816(c[mt][nt], a[mt][1], bf[nt][2], bf[nt][3]);
      }
    __syncwarp();                                     // every lane is done reading sf / so(pairs) as operands
    // dF goes out through the feature tile itself (it has been consumed); row 0 adds the dense passthrough
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int i = mt * 16 + g + h * 8, d = nt * 8 + tq * 2;
          if (i < S) {
            float v0 = c[mt][nt][h * 2], v1 = c[mt][nt][h * 2 + 1];
            if (i == 0) { v0 += __bfloat162float(so[d]); v1 += __bfloat162float(so[d + 1]); }
            *reinterpret_cast<uint32_t*>(&sf[i][d]) = pack_bf16x2(v0, v1);
          }
        }
    __syncwarp();
    uint4* df = reinterpret_cast<uint4*>(dfeats + size_t(b) * S * kDiD);
    for (int cc = lane; cc < S * 4; cc += 32) df[cc] = *reinterpret_cast<const uint4*>(&sf[cc >> 2][(cc & 3) * 8]);
    __syncwarp();
  }
}

bool dot_interaction_mma_ok(int S, int D, int ld_out, const void* a, const void* b, const void* c) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  return D == kDiD && S >= 2 && S <= 32 && ld_out % 8 == 0 && ld_out <= kDiMaxOut &&
         ld_out >= kDiD + S * (S - 1) / 2 && al(a) && al(b) && al(c);
}

int dot_interaction_fwd_mma(const void* feats, void* out, int B, int S, int ld_out, cudaStream_t stream) {
  int grid = (B + kDiWarps - 1) / kDiWarps;
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  dot_interaction_fwd_mma_kernel<<<grid, kDiWarps * 32, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(feats), static_cast<__nv_bfloat16*>(out), B, S, ld_out);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int dot_interaction_bwd_mma(const void* feats, const void* dout, void* dfeats, int B, int S, int ld_out,
                            cudaStream_t stream) {
  int grid = (B + kDiWarps - 1) / kDiWarps;
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  dot_interaction_bwd_mma_kernel<<<grid, kDiWarps * 32, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(feats), static_cast<const __nv_bfloat16*>(dout),
      static_cast<__nv_bfloat16*>(dfeats), B, S, ld_out);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace rdp
