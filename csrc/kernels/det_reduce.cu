// Ordered (bit-reproducible) column sums of bf16 matrices: the deterministic route for bias gradients.
// The fused epilogues add their per-CTA column sums with fp32 atomics, whose arrival order -- and with it the
// last bits of the sum -- changes from run to run; here every column is summed by a fixed tree instead:
// the rows are cut into R chunks (one CTA each); inside a CTA 8 warps take rows w, w + 8, ... sequentially and
// their 8 partial sums are added in warp order; the last CTA of a column group to finish (whichever it is)
// adds the R chunk sums in chunk order.  Several matrices (all layers of a step) are handled by one launch.
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

constexpr int kColsumMaxSeg = 16;
constexpr int kColsumMaxChunks = 32;

struct ColsumParams {
  const __nv_bfloat16* src[kColsumMaxSeg];
  float* out[kColsumMaxSeg];
  int ld[kColsumMaxSeg];
  int n[kColsumMaxSeg];           // columns to sum (out[0..n))
  int first_group[kColsumMaxSeg + 1];
  int n_seg, rows, rows_per_chunk, groups;
  float* ws;                      // [chunks, groups, 64] chunk sums
  unsigned int* counters;         // [groups], zero between launches
};

__global__ void __launch_bounds__(256)
colsum_ordered_kernel(const __grid_constant__ ColsumParams p) {
  __shared__ float s_part[8][64];
  __shared__ bool s_last;
  const int group = blockIdx.x, chunk = blockIdx.y, chunks = gridDim.y;
  int seg = 0;
  while (seg + 1 < p.n_seg && group >= p.first_group[seg + 1]) ++seg;
  const int c0 = (group - p.first_group[seg]) * 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = c0 + lane * 2;
  const int r_end = min(p.rows, (chunk + 1) * p.rows_per_chunk);
  float a0 = 0.f, a1 = 0.f;
  if (c < p.n[seg]) {
    const int ld = p.ld[seg];
    const __nv_bfloat16* src = p.src[seg] + c;
    const bool pair = c + 1 < p.n[seg] && (ld & 1) == 0;
    int r = chunk * p.rows_per_chunk + warp;
    if (pair) {
      for (; r + 24 < r_end; r += 32) {           // four independent loads in flight; added in row order
        const uint32_t v0 = *reinterpret_cast<const uint32_t*>(src + size_t(r) * ld);
        const uint32_t v1 = *reinterpret_cast<const uint32_t*>(src + size_t(r + 8) * ld);
        const uint32_t v2 = *reinterpret_cast<const uint32_t*>(src + size_t(r + 16) * ld);
        const uint32_t v3 = *reinterpret_cast<const uint32_t*>(src + size_t(r + 24) * ld);
        float2 f = unpack_bf16x2(v0); a0 += f.x; a1 += f.y;
        f = unpack_bf16x2(v1); a0 += f.x; a1 += f.y;
        f = unpack_bf16x2(v2); a0 += f.x; a1 += f.y;
        f = unpack_bf16x2(v3); a0 += f.x; a1 += f.y;
      }
    }
    for (; r < r_end; r += 8) {
      if (pair) {
        const float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(src + size_t(r) * ld));
        a0 += v.x; a1 += v.y;
      } else {
        a0 += __bfloat162float(src[size_t(r) * ld]);
        if (c + 1 < p.n[seg]) a1 += __bfloat162float(src[size_t(r) * ld + 1]);
      }
    }
  }
  s_part[warp][lane * 2] = a0;
  s_part[warp][lane * 2 + 1] = a1;
  __syncthreads();
  if (threadIdx.x < 64) {
    float sum = s_part[0][threadIdx.x];
#pragma unroll
    for (int w = 1; w < 8; ++w) sum += s_part[w][threadIdx.x];
    if (chunks == 1) {
      if (c0 + int(threadIdx.x) < p.n[seg]) p.out[seg][c0 + threadIdx.x] = sum;
    } else {
      p.ws[(size_t(chunk) * p.groups + group) * 64 + threadIdx.x] = sum;
    }
  }
  if (chunks == 1) return;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(p.counters + group, 1u);
    s_last = prev == unsigned(chunks - 1);
    if (s_last) p.counters[group] = 0;
  }
  __syncthreads();
  if (s_last && threadIdx.x < 64 && c0 + int(threadIdx.x) < p.n[seg]) {
    __threadfence();
    float sum = 0.f;
    for (int k = 0; k < chunks; ++k)
      sum += *reinterpret_cast<volatile float*>(p.ws + (size_t(k) * p.groups + group) * 64 + threadIdx.x);
    p.out[seg][c0 + threadIdx.x] = sum;
  }
}

}  // namespace rdp

using namespace rdp;

// ws: fp32 scratch of ws_floats values followed by nothing else; the first `groups` 32-bit words of `counters`
// must be zero (the kernel leaves them zero).  The number of row chunks adapts to the scratch size.
extern "C" int rdp_colsum_ordered(const void* const* src, const int* ld, const int* n, float* const* out, int n_seg,
                                  int rows, float* ws, long long ws_floats, unsigned int* counters, int n_counters,
                                  void* stream) {
  RDP_REQUIRE(n_seg >= 1 && n_seg <= kColsumMaxSeg, "rdp_colsum_ordered: 1..%d matrices per call (got %d)",
              kColsumMaxSeg, n_seg);
  if (rows <= 0) return 0;
  ColsumParams p;
  memset(&p, 0, sizeof(p));
  int groups = 0;
  for (int i = 0; i < n_seg; ++i) {
    p.src[i] = static_cast<const __nv_bfloat16*>(src[i]); p.out[i] = out[i]; p.ld[i] = ld[i]; p.n[i] = n[i];
    p.first_group[i] = groups;
    groups += (n[i] + 63) / 64;
  }
  p.first_group[n_seg] = groups;
  if (groups == 0) return 0;
  int chunks = (rows + 255) / 256;
  if (chunks > kColsumMaxChunks) chunks = kColsumMaxChunks;
  if (ws == nullptr || counters == nullptr || n_counters < groups) chunks = 1;
  else if ((long long)chunks * groups * 64 > ws_floats) chunks = int(ws_floats / (groups * 64LL));
  if (chunks < 1) chunks = 1;
  int rpc = (rows + chunks - 1) / chunks;
  rpc = (rpc + 7) / 8 * 8;
  chunks = (rows + rpc - 1) / rpc;
  p.n_seg = n_seg; p.rows = rows; p.rows_per_chunk = rpc; p.groups = groups; p.ws = ws; p.counters = counters;
  colsum_ordered_kernel<<<dim3(groups, chunks), 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
