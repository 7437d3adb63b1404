// Histogram gradient boosting on the GPU: the two per-level passes over the rows of `XGBoostEstimator`'s tree
// builder (reference: python/raydp/xgboost/estimator.py hands the data to xgboost-ray; here the booster is ours,
// raydp_b200/xgboost/hist_gbt.py, and these kernels are its `tree_method="gpu_hist"`).
//
//   gbt_hist_kernel       for every row of the level and every feature: hist[slot][g|h][feature][bin] += g|h.
//                         Rows are stored FEATURE-MAJOR (binned_t[f][row], uint16), so a warp reads 64 contiguous
//                         bytes per feature.  A block owns one feature and a chunk of rows and builds its histogram
//                         in SHARED memory (fp32, shared atomics; at most a few thousand rows per block, so fp32
//                         partial sums are exact enough), then adds the non-empty bins to the global fp64 histogram:
//                         global atomics drop from 2 per (row, feature) to 2 per touched (block, slot, bin).
//                         Levels too wide for shared memory (slots * bins * 8 B > 200 KB) add to global directly.
//   gbt_partition_kernel  after the splits of a level are chosen: every row moves to its child node and gets
//                         its histogram slot in the next level.
// The small [slots, 2, F, bins] histogram goes back to the host, where the split search runs in fp64 and the
// worker group all-reduces it (one collective per level).
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

struct GbtHistParams {
  const uint16_t* binned_t;   // [F, n]
  const float* g;             // [n]
  const float* h;             // [n]
  const int* slot;            // [n] slot of the row's node in this level, -1 = row not in the level / not sampled
  double* hist;               // [n_slots, 2, F, n_bins], zeroed by the caller
  long long n;
  int F, n_bins, n_slots;
  int rows_per_block;
  int use_smem;
};

__global__ void __launch_bounds__(256)
gbt_hist_kernel(const GbtHistParams p) {
  extern __shared__ float sh[];                       // [n_slots][n_bins][2] when use_smem
  const int f = blockIdx.x;
  const long long r0 = (long long)blockIdx.y * p.rows_per_block;
  const long long r1 = min(p.n, r0 + p.rows_per_block);
  const int cells = p.n_slots * p.n_bins;
  const uint16_t* col = p.binned_t + (size_t)f * p.n;
  if (p.use_smem) {
    for (int i = threadIdx.x; i < 2 * cells; i += blockDim.x) sh[i] = 0.f;
    __syncthreads();
    for (long long r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
      const int s = p.slot[r];
      if (s < 0) continue;
      const int cell = s * p.n_bins + col[r];
      atomicAdd(&sh[2 * cell], p.g[r]);
      atomicAdd(&sh[2 * cell + 1], p.h[r]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < cells; i += blockDim.x) {
      const float gs = sh[2 * i], hs = sh[2 * i + 1];
      if (gs == 0.f && hs == 0.f) continue;
      const int s = i / p.n_bins, b = i % p.n_bins;
      double* base = p.hist + ((size_t)s * 2 * p.F + f) * p.n_bins + b;
      atomicAdd(base, (double)gs);
      atomicAdd(base + (size_t)p.F * p.n_bins, (double)hs);
    }
  } else {
    for (long long r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
      const int s = p.slot[r];
      if (s < 0) continue;
      double* base = p.hist + ((size_t)s * 2 * p.F + f) * p.n_bins + col[r];
      atomicAdd(base, (double)p.g[r]);
      atomicAdd(base + (size_t)p.F * p.n_bins, (double)p.h[r]);
    }
  }
}

struct GbtPartitionParams {
  const uint16_t* binned_t;   // [F, n]
  int* node_of;               // [n] in / out: tree node of every row
  int* slot;                  // [n] out: histogram slot in the NEXT level (-1: not sampled or node is a leaf)
  const uint8_t* in_sample;   // [n] row takes part in this tree's histograms
  const int* feat;            // [nodes] split feature, -1 = not split
  const int* thr_bin;         // [nodes] rows with bin <= thr_bin go left
  const int* left;            // [nodes]
  const int* right;           // [nodes]
  const uint8_t* default_left;  // [nodes] where rows with a missing value go
  const int* next_slot;       // [nodes] slot in the next level, -1 = none
  long long n;
  int miss_bin;
};

__global__ void __launch_bounds__(256)
gbt_partition_kernel(const GbtPartitionParams p) {
  for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < p.n; r += (long long)gridDim.x * blockDim.x) {
    int nid = p.node_of[r];
    const int f = p.feat[nid];
    if (f >= 0) {
      const int b = p.binned_t[(size_t)f * p.n + r];
      const bool go_left = (b == p.miss_bin) ? (p.default_left[nid] != 0) : (b <= p.thr_bin[nid]);
      nid = go_left ? p.left[nid] : p.right[nid];
      p.node_of[r] = nid;
    }
    p.slot[r] = p.in_sample[r] ? p.next_slot[nid] : -1;
  }
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_gbt_hist(const void* binned_t, const float* g, const float* h, const int* slot, double* hist, long long n,
                 int F, int n_bins, int n_slots, void* stream) {
  RDP_REQUIRE(n >= 0 && F >= 1 && n_bins >= 2 && n_slots >= 1, "rdp_gbt_hist: bad shape n=%lld F=%d bins=%d slots=%d",
              n, F, n_bins, n_slots);
  if (n == 0) return 0;
  GbtHistParams p;
  p.binned_t = static_cast<const uint16_t*>(binned_t); p.g = g; p.h = h; p.slot = slot; p.hist = hist;
  p.n = n; p.F = F; p.n_bins = n_bins; p.n_slots = n_slots;
  const size_t smem = (size_t)n_slots * n_bins * 2 * sizeof(float);
  p.use_smem = smem <= 200 * 1024 ? 1 : 0;
  // enough blocks to fill the machine, few enough rows per block for fp32 partial sums (<= 16384)
  long long chunks = (long long)num_sms() * 8 / F;
  if (chunks < 1) chunks = 1;
  long long rpb = (n + chunks - 1) / chunks;
  if (rpb > 16384) rpb = 16384;
  if (rpb < 1024) rpb = 1024;
  p.rows_per_block = int(rpb);
  chunks = (n + rpb - 1) / rpb;
  RDP_REQUIRE(chunks <= 65535, "rdp_gbt_hist: %lld row chunks (more than 1 G rows per call?)", chunks);
  if (p.use_smem) RDP_SMEM_ATTR_ONCE_PER_DEVICE(gbt_hist_kernel, 227 * 1024);
  gbt_hist_kernel<<<dim3(F, (unsigned)chunks), 256, p.use_smem ? smem : 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_gbt_partition(const void* binned_t, int* node_of, int* slot, const void* in_sample, const int* feat,
                      const int* thr_bin, const int* left, const int* right, const void* default_left,
                      const int* next_slot, long long n, int miss_bin, void* stream) {
  if (n <= 0) return 0;
  GbtPartitionParams p;
  p.binned_t = static_cast<const uint16_t*>(binned_t); p.node_of = node_of; p.slot = slot;
  p.in_sample = static_cast<const uint8_t*>(in_sample); p.feat = feat; p.thr_bin = thr_bin; p.left = left;
  p.right = right; p.default_left = static_cast<const uint8_t*>(default_left); p.next_slot = next_slot;
  p.n = n; p.miss_bin = miss_bin;
  long long grid = (n + 255) / 256;
  if (grid > (long long)num_sms() * 16) grid = (long long)num_sms() * 16;
  gbt_partition_kernel<<<(unsigned)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
