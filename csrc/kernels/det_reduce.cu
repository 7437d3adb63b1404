// Ordered (bit-reproducible) column sums of bf16 matrices: the deterministic route for bias gradients.
// The fused epilogues add their per-CTA column sums with fp32 atomics, whose arrival order -- and with it the
// last bits of the sum -- changes from run to run; here every column is summed by a fixed tree instead:
// the rows are cut into R chunks (one CTA per chunk and 128 columns); inside a CTA 8 warps take rows w, w + 8, ... sequentially and
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
  float* ws;                      // [chunks, groups, 128] chunk sums
  unsigned int* counters;         // [groups], zero between launches
};

constexpr int kColsumCols = 128;    // columns per CTA: 4 per lane, read as one 8-byte load

__global__ void __launch_bounds__(256)
colsum_ordered_kernel(const __grid_constant__ ColsumParams p) {
  __shared__ float s_part[8][kColsumCols];
  __shared__ bool s_last;
  const int group = blockIdx.x, chunk = blockIdx.y, chunks = gridDim.y;
  int seg = 0;
  while (seg + 1 < p.n_seg && group >= p.first_group[seg + 1]) ++seg;
  const int c0 = (group - p.first_group[seg]) * kColsumCols;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = c0 + lane * 4;
  const int r_end = min(p.rows, (chunk + 1) * p.rows_per_chunk);
  const int ncol = p.n[seg];
  float a[4] = {0.f, 0.f, 0.f, 0.f};
  if (c < ncol) {
    const int ld = p.ld[seg];
    const __nv_bfloat16* src = p.src[seg] + c;
    const bool quad = c + 3 < ncol && (ld & 3) == 0 && (reinterpret_cast<uintptr_t>(p.src[seg]) & 7) == 0;
    int r = chunk * p.rows_per_chunk + warp;
    if (quad) {
      for (; r + 24 < r_end; r += 32) {           // four independent loads in flight; added in row order
        uint2 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const uint2*>(src + size_t(r + 8 * u) * ld);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float2 lo = unpack_bf16x2(v[u].x), hi = unpack_bf16x2(v[u].y);
          a[0] += lo.x; a[1] += lo.y; a[2] += hi.x; a[3] += hi.y;
        }
      }
      for (; r < r_end; r += 8) {
        const uint2 v = *reinterpret_cast<const uint2*>(src + size_t(r) * ld);
        const float2 lo = unpack_bf16x2(v.x), hi = unpack_bf16x2(v.y);
        a[0] += lo.x; a[1] += lo.y; a[2] += hi.x; a[3] += hi.y;
      }
    } else {
      for (; r < r_end; r += 8) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (c + k < ncol) a[k] += __bfloat162float(src[size_t(r) * ld + k]);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) s_part[warp][lane * 4 + k] = a[k];
  __syncthreads();
  if (threadIdx.x < kColsumCols) {
    float sum = s_part[0][threadIdx.x];
#pragma unroll
    for (int w = 1; w < 8; ++w) sum += s_part[w][threadIdx.x];
    if (chunks == 1) {
      if (c0 + int(threadIdx.x) < ncol) p.out[seg][c0 + threadIdx.x] = sum;
    } else {
      p.ws[(size_t(chunk) * p.groups + group) * kColsumCols + threadIdx.x] = sum;
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
  if (s_last && threadIdx.x < kColsumCols && c0 + int(threadIdx.x) < ncol) {
    __threadfence();
    float sum = 0.f;
    for (int k = 0; k < chunks; ++k)
      sum += *reinterpret_cast<volatile float*>(p.ws + (size_t(k) * p.groups + group) * kColsumCols + threadIdx.x);
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
    groups += (n[i] + kColsumCols - 1) / kColsumCols;
  }
  p.first_group[n_seg] = groups;
  if (groups == 0) return 0;
  int chunks = (rows + 255) / 256;
  if (chunks > kColsumMaxChunks) chunks = kColsumMaxChunks;
  if (ws == nullptr || counters == nullptr || n_counters < groups) chunks = 1;
  else if ((long long)chunks * groups * kColsumCols > ws_floats) chunks = int(ws_floats / (groups * (long long)kColsumCols));
  if (chunks < 1) chunks = 1;
  int rpc = (rows + chunks - 1) / chunks;
  rpc = (rpc + 7) / 8 * 8;
  chunks = (rows + rpc - 1) / rpc;
  p.n_seg = n_seg; p.rows = rows; p.rows_per_chunk = rpc; p.groups = groups; p.ws = ws; p.counters = counters;
  colsum_ordered_kernel<<<dim3(groups, chunks), 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
