// K5: ingestion -- Arrow column buffers (already DMA'd to the device, column-major, mixed
// dtypes) -> the training tensors, in one pass.
//
// Replaces reference python/raydp/torch/estimator.py:227-232 (`dataset.to_torch`: Arrow ->
// pandas -> numpy -> torch.as_tensor -> concat + cast on the CPU).  The host never touches the
// values: pinned Arrow buffers are copied as they are (cudaMemcpyAsync on a side stream) and
// this kernel decodes f32/f64/i32/i64/i16/i8 columns, applies an optional per-column
// (x - shift) * scale normalisation, writes the zero-padded row-major [rows, ld] bf16 (or f32)
// feature matrix the tcgen05 GEMM reads through TMA, and the f32 label vector.  For DLRM-style
// inputs rdp_ingest_indices turns categorical columns into int32 table rows (modulo the table
// size) and can append hashed feature crosses of column pairs.
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

constexpr int kMaxCols = 64;

struct IngestParams {
  const void* col[kMaxCols];
  uint8_t dtype[kMaxCols];      // 0 f32, 1 f64, 2 i32, 3 i64, 4 i16, 5 i8/u8
  int n_cols;
  int64_t rows;
  int64_t src_offset;           // first row inside the column buffers
  void* out;                    // [rows, ld] bf16 or f32
  int ld;
  int out_f32;
  const float* shift;           // [n_cols] or null
  const float* scale;           // [n_cols] or null
  const void* label;            // label column (same dtype codes) or null
  uint8_t label_dtype;
  float* label_out;             // [rows]
};

__device__ __forceinline__ float load_as_float(const void* base, uint8_t dt, int64_t i) {
  switch (dt) {
    case 0: return __ldg(static_cast<const float*>(base) + i);
    case 1: return float(__ldg(static_cast<const double*>(base) + i));
    case 2: return float(__ldg(static_cast<const int*>(base) + i));
    case 3: return float(__ldg(static_cast<const long long*>(base) + i));
    case 4: return float(__ldg(static_cast<const short*>(base) + i));
    default: return float(__ldg(static_cast<const signed char*>(base) + i));
  }
}

__global__ void __launch_bounds__(256)
ingest_pack_kernel(const IngestParams p) {
  // one thread per row: column reads are coalesced across the warp (same column, consecutive
  // rows); each thread emits its row as 16-byte vectors.
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < p.rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int64_t src = r + p.src_offset;
    for (int c0 = 0; c0 < p.ld; c0 += 8) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int c = c0 + j;
        float x = 0.f;
        if (c < p.n_cols) {
          x = load_as_float(p.col[c], p.dtype[c], src);
          if (p.shift) x -= p.shift[c];
          if (p.scale) x *= p.scale[c];
        }
        v[j] = x;
      }
      if (p.out_f32) {
        float4* o = reinterpret_cast<float4*>(static_cast<float*>(p.out) + r * p.ld + c0);
        o[0] = make_float4(v[0], v[1], v[2], v[3]);
        o[1] = make_float4(v[4], v[5], v[6], v[7]);
      } else {
        uint4* o = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + r * p.ld + c0);
        *o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                        pack_bf16x2(v[6], v[7]));
      }
    }
    if (p.label != nullptr) p.label_out[r] = load_as_float(p.label, p.label_dtype, src);
  }
}

struct IndexParams {
  const void* col[kMaxCols];
  uint8_t dtype[kMaxCols];      // 2 i32, 3 i64
  int n_cols;
  int64_t rows;
  int64_t src_offset;
  int32_t table_rows[kMaxCols]; // modulo per output column (0 = no modulo)
  int n_cross;                  // hashed crosses appended after the plain columns
  uint8_t cross_a[16], cross_b[16];
  int32_t cross_rows[16];
  int* out;                     // [rows, n_cols + n_cross] int32
};

__device__ __forceinline__ long long load_as_i64(const void* base, uint8_t dt, int64_t i) {
  return dt == 3 ? __ldg(static_cast<const long long*>(base) + i)
                 : (long long)__ldg(static_cast<const int*>(base) + i);
}

__device__ __forceinline__ uint32_t mix64(unsigned long long a, unsigned long long b) {
  unsigned long long h = (a * 0x9E3779B97F4A7C15ull) ^ (b + 0x7F4A7C15ull + (a << 6) + (a >> 2));
  h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
  return uint32_t(h);
}

__global__ void __launch_bounds__(256)
ingest_indices_kernel(const IndexParams p) {
  const int width = p.n_cols + p.n_cross;
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < p.rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int64_t src = r + p.src_offset;
    int* o = p.out + r * width;
    for (int c = 0; c < p.n_cols; ++c) {
      long long x = load_as_i64(p.col[c], p.dtype[c], src);
      if (p.table_rows[c] > 0) { x %= p.table_rows[c]; if (x < 0) x += p.table_rows[c]; }
      o[c] = int(x);
    }
    for (int k = 0; k < p.n_cross; ++k) {
      const long long a = load_as_i64(p.col[p.cross_a[k]], p.dtype[p.cross_a[k]], src);
      const long long b = load_as_i64(p.col[p.cross_b[k]], p.dtype[p.cross_b[k]], src);
      o[p.n_cols + k] = int(mix64((unsigned long long)a, (unsigned long long)b) % uint32_t(p.cross_rows[k]));
    }
  }
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_ingest_pack(const void* const* cols, const int* dtypes, int n_cols, int64_t rows, int64_t src_offset,
                    void* out, int ld, int out_f32, const float* shift, const float* scale,
                    const void* label, int label_dtype, float* label_out, void* stream) {
  RDP_REQUIRE(n_cols >= 0 && n_cols <= kMaxCols, "rdp_ingest_pack: at most %d feature columns (got %d)", kMaxCols, n_cols);
  RDP_REQUIRE(ld % 8 == 0 && ld >= n_cols, "rdp_ingest_pack: ld=%d must be a multiple of 8 and >= n_cols=%d", ld, n_cols);
  if (rows <= 0) return 0;
  IngestParams p;
  for (int i = 0; i < n_cols; ++i) { p.col[i] = cols[i]; p.dtype[i] = uint8_t(dtypes[i]); }
  p.n_cols = n_cols; p.rows = rows; p.src_offset = src_offset; p.out = out; p.ld = ld; p.out_f32 = out_f32;
  p.shift = shift; p.scale = scale; p.label = label; p.label_dtype = uint8_t(label_dtype); p.label_out = label_out;
  int grid = int((rows + 255) / 256);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  ingest_pack_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int rdp_ingest_indices(const void* const* cols, const int* dtypes, int n_cols, int64_t rows, int64_t src_offset,
                       const int* table_rows, int n_cross, const int* cross_a, const int* cross_b,
                       const int* cross_rows, int* out, void* stream) {
  RDP_REQUIRE(n_cols > 0 && n_cols <= kMaxCols && n_cross >= 0 && n_cross <= 16,
              "rdp_ingest_indices: bad column counts %d / %d", n_cols, n_cross);
  if (rows <= 0) return 0;
  IndexParams p;
  for (int i = 0; i < n_cols; ++i) {
    RDP_REQUIRE(dtypes[i] == 2 || dtypes[i] == 3, "rdp_ingest_indices: column %d is not int32/int64", i);
    p.col[i] = cols[i]; p.dtype[i] = uint8_t(dtypes[i]); p.table_rows[i] = table_rows ? table_rows[i] : 0;
  }
  p.n_cols = n_cols; p.rows = rows; p.src_offset = src_offset; p.n_cross = n_cross;
  for (int k = 0; k < n_cross; ++k) {
    p.cross_a[k] = uint8_t(cross_a[k]); p.cross_b[k] = uint8_t(cross_b[k]); p.cross_rows[k] = cross_rows[k];
  }
  p.out = out;
  int grid = int((rows + 255) / 256);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  ingest_indices_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
