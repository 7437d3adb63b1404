// K5: ingestion -- Arrow column buffers (already DMA'd to the device, column-major, mixed
// dtypes) -> the training tensors, in one pass.
//
// Replaces reference python/raydp/torch/estimator.py:227-232 (`dataset.to_torch`: Arrow ->
// pandas -> numpy -> torch.as_tensor -> concat + cast on the CPU).  The host never touches the
// values: pinned Arrow buffers are copied as they are (rdp_ingest_stage: one call issues every
// cudaMemcpyAsync of a chunk on a side stream) and the pack kernel decodes them on the device:
//
//   * f32/f64/i32/i64/i16/i8 value buffers;
//   * Arrow validity bitmaps (bit i of the bitmap = row i is valid): a null becomes `null_value`
//     (NaN by default on the feature path -- what the reference's pandas/numpy path yields for a null
//     float -- or any fill value the caller asks for); nulls in index columns become row 0;
//   * vector columns (fixed_size_list<T>[w] / list<T> of one length, e.g. a VectorAssembler output):
//     the child value buffer is row-major [rows, w] and contributes w consecutive features;
//   * an optional per-feature (x - shift) * scale normalisation;
//
// and writes the zero-padded row-major [rows, ld] bf16 (or f32) feature matrix the tcgen05 GEMM
// reads through TMA, plus the f32 label vector.  For DLRM-style inputs rdp_ingest_indices turns
// categorical columns into int32 table rows (modulo the table size) and can append hashed feature
// crosses of column pairs.
#include "common.cuh"
#include "raydp_b200.h"

namespace rdp {

constexpr int kMaxCols = 64;

struct IngestParams {
  const void* col[kMaxCols];    // value buffer of source column s (row-major [rows, width[s]])
  const uint8_t* valid[kMaxCols];   // validity bitmap of source column s or null (all valid)
  int32_t vbit0[kMaxCols];      // bit index of row 0 of the staged range inside valid[s]
  int32_t width[kMaxCols];      // 1 for scalar columns, w for vector columns
  uint8_t dtype[kMaxCols];      // 0 f32, 1 f64, 2 i32, 3 i64, 4 i16, 5 i8/u8
  int n_src;                    // source columns
  int n_feat;                   // sum of widths
  int64_t rows;
  int64_t src_offset;           // first row inside the column buffers
  void* out;                    // [rows, ld] bf16 or f32
  int ld;
  int out_f32;
  const float* shift;           // [n_feat] or null
  const float* scale;           // [n_feat] or null
  float null_value;
  const void* label;            // label column (same dtype codes) or null
  const uint8_t* label_valid;
  int32_t label_vbit0;
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

__device__ __forceinline__ bool bit_set(const uint8_t* bits, int64_t i) {
  return (__ldg(bits + (i >> 3)) >> (i & 7)) & 1;
}

__global__ void __launch_bounds__(256)
ingest_pack_kernel(const __grid_constant__ IngestParams p) {
  // one thread per row: column reads are coalesced across the warp (same column, consecutive
  // rows); each thread emits its row as 16-byte vectors.  The walk over (source column, inner
  // element) is the same for every thread, so the loop below never diverges.
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < p.rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int64_t src = r + p.src_offset;
    int s = 0, inner = 0;
    bool ok = p.n_src > 0 && (p.valid[0] == nullptr || bit_set(p.valid[0], src + p.vbit0[0]));
    for (int c0 = 0; c0 < p.ld; c0 += 8) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int c = c0 + j;
        float x = 0.f;
        if (c < p.n_feat) {
          const int w = p.width[s];
          x = ok ? load_as_float(p.col[s], p.dtype[s], src * w + inner) : p.null_value;
          if (ok || p.null_value == p.null_value) {       // a NaN fill stays NaN; finite fills are normalised
            if (p.shift) x -= p.shift[c];
            if (p.scale) x *= p.scale[c];
          }
          if (++inner == w) {
            inner = 0;
            ++s;
            ok = s < p.n_src && (p.valid[s] == nullptr || bit_set(p.valid[s], src + p.vbit0[s]));
          }
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
    if (p.label != nullptr) {
      const bool lok = p.label_valid == nullptr || bit_set(p.label_valid, src + p.label_vbit0);
      p.label_out[r] = lok ? load_as_float(p.label, p.label_dtype, src) : p.null_value;
    }
  }
}

struct IndexParams {
  const void* col[kMaxCols];
  const uint8_t* valid[kMaxCols];
  int32_t vbit0[kMaxCols];
  uint8_t dtype[kMaxCols];      // 2 i32, 3 i64, 4 i16, 5 i8
  int n_cols;
  int64_t rows;
  int64_t src_offset;
  int32_t table_rows[kMaxCols]; // > 0: modulo (negatives wrap, null -> row 0); 0: as is; < 0: multi-hot bag column of a
                                // table with -table_rows rows: negative or null -> -1 (padding), else modulo
  int n_cross;                  // hashed crosses appended after the plain columns
  uint8_t cross_a[16], cross_b[16];
  int32_t cross_rows[16];
  int* out;                     // [rows, n_cols + n_cross] int32
};

__device__ __forceinline__ long long load_as_i64(const void* base, uint8_t dt, int64_t i) {
  switch (dt) {
    case 3: return __ldg(static_cast<const long long*>(base) + i);
    case 2: return (long long)__ldg(static_cast<const int*>(base) + i);
    case 4: return (long long)__ldg(static_cast<const short*>(base) + i);
    default: return (long long)__ldg(static_cast<const signed char*>(base) + i);
  }
}

__device__ __forceinline__ uint32_t mix64(unsigned long long a, unsigned long long b) {
  unsigned long long h = (a * 0x9E3779B97F4A7C15ull) ^ (b + 0x7F4A7C15ull + (a << 6) + (a >> 2));
  h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
  return uint32_t(h);
}

__device__ __forceinline__ long long index_value(const IndexParams& p, int c, int64_t src) {
  if (p.valid[c] != nullptr && !bit_set(p.valid[c], src + p.vbit0[c])) return 0;    // null -> row 0
  return load_as_i64(p.col[c], p.dtype[c], src);
}

__global__ void __launch_bounds__(256)
ingest_indices_kernel(const __grid_constant__ IndexParams p) {
  const int width = p.n_cols + p.n_cross;
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < p.rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int64_t src = r + p.src_offset;
    int* o = p.out + r * width;
    for (int c = 0; c < p.n_cols; ++c) {
      long long x = index_value(p, c, src);
      const int tr = p.table_rows[c];
      if (tr > 0) { x %= tr; if (x < 0) x += tr; }
      else if (tr < 0) {
        const bool null = p.valid[c] != nullptr && !bit_set(p.valid[c], src + p.vbit0[c]);
        x = (null || x < 0) ? -1 : x % (-tr);
      }
      o[c] = int(x);
    }
    for (int k = 0; k < p.n_cross; ++k) {
      const long long a = index_value(p, p.cross_a[k], src);
      const long long b = index_value(p, p.cross_b[k], src);
      o[p.n_cols + k] = int(mix64((unsigned long long)a, (unsigned long long)b) % uint32_t(p.cross_rows[k]));
    }
  }
}

static int launch_pack(const IngestParams& p, cudaStream_t stream) {
  int grid = int((p.rows + 255) / 256);
  if (grid > num_sms() * 8) grid = num_sms() * 8;
  ingest_pack_kernel<<<grid, 256, 0, stream>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
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
  for (int i = 0; i < n_cols; ++i) {
    p.col[i] = cols[i]; p.dtype[i] = uint8_t(dtypes[i]); p.valid[i] = nullptr; p.vbit0[i] = 0; p.width[i] = 1;
  }
  p.n_src = n_cols; p.n_feat = n_cols; p.rows = rows; p.src_offset = src_offset; p.out = out; p.ld = ld;
  p.out_f32 = out_f32; p.shift = shift; p.scale = scale; p.null_value = 0.f;
  p.label = label; p.label_valid = nullptr; p.label_vbit0 = 0; p.label_dtype = uint8_t(label_dtype);
  p.label_out = label_out;
  return launch_pack(p, static_cast<cudaStream_t>(stream));
}

int rdp_ingest_pack_v2(const void* const* cols, const void* const* valid, const int* vbit0, const int* widths,
                       const int* dtypes, int n_src, int64_t rows, int64_t src_offset, void* out, int ld,
                       int out_f32, const float* shift, const float* scale, float null_value, const void* label,
                       const void* label_valid, int label_vbit0, int label_dtype, float* label_out, void* stream) {
  RDP_REQUIRE(n_src >= 0 && n_src <= kMaxCols, "rdp_ingest_pack_v2: at most %d source columns (got %d)", kMaxCols, n_src);
  if (rows <= 0) return 0;
  IngestParams p;
  int n_feat = 0;
  for (int i = 0; i < n_src; ++i) {
    RDP_REQUIRE(widths[i] >= 1, "rdp_ingest_pack_v2: column %d has width %d", i, widths[i]);
    p.col[i] = cols[i]; p.dtype[i] = uint8_t(dtypes[i]);
    p.valid[i] = valid ? static_cast<const uint8_t*>(valid[i]) : nullptr;
    p.vbit0[i] = vbit0 ? vbit0[i] : 0; p.width[i] = widths[i];
    n_feat += widths[i];
  }
  RDP_REQUIRE(ld % 8 == 0 && ld >= n_feat, "rdp_ingest_pack_v2: ld=%d must be a multiple of 8 and >= %d features", ld, n_feat);
  p.n_src = n_src; p.n_feat = n_feat; p.rows = rows; p.src_offset = src_offset; p.out = out; p.ld = ld;
  p.out_f32 = out_f32; p.shift = shift; p.scale = scale; p.null_value = null_value;
  p.label = label; p.label_valid = static_cast<const uint8_t*>(label_valid); p.label_vbit0 = label_vbit0;
  p.label_dtype = uint8_t(label_dtype); p.label_out = label_out;
  return launch_pack(p, static_cast<cudaStream_t>(stream));
}

int rdp_ingest_indices(const void* const* cols, const int* dtypes, int n_cols, int64_t rows, int64_t src_offset,
                       const int* table_rows, int n_cross, const int* cross_a, const int* cross_b,
                       const int* cross_rows, int* out, void* stream) {
  return rdp_ingest_indices_v2(cols, nullptr, nullptr, dtypes, n_cols, rows, src_offset, table_rows, n_cross,
                               cross_a, cross_b, cross_rows, out, stream);
}

int rdp_ingest_indices_v2(const void* const* cols, const void* const* valid, const int* vbit0, const int* dtypes,
                          int n_cols, int64_t rows, int64_t src_offset, const int* table_rows, int n_cross,
                          const int* cross_a, const int* cross_b, const int* cross_rows, int* out, void* stream) {
  RDP_REQUIRE(n_cols > 0 && n_cols <= kMaxCols && n_cross >= 0 && n_cross <= 16,
              "rdp_ingest_indices: bad column counts %d / %d", n_cols, n_cross);
  if (rows <= 0) return 0;
  IndexParams p;
  for (int i = 0; i < n_cols; ++i) {
    RDP_REQUIRE(dtypes[i] >= 2 && dtypes[i] <= 5, "rdp_ingest_indices: column %d is not an integer column", i);
    p.col[i] = cols[i]; p.dtype[i] = uint8_t(dtypes[i]); p.table_rows[i] = table_rows ? table_rows[i] : 0;
    p.valid[i] = valid ? static_cast<const uint8_t*>(valid[i]) : nullptr;
    p.vbit0[i] = vbit0 ? vbit0[i] : 0;
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

int rdp_ingest_stage(const uint64_t* dst, const uint64_t* src_host, const uint64_t* bytes, int n, void* stream) {
  // every host->device copy of one chunk from ONE native call: a plan array in, one cudaMemcpyAsync per
  // span inside (the spans are the Arrow buffers as they lie in pinned shared memory)
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  for (int i = 0; i < n; ++i) {
    if (bytes[i] == 0) continue;
    RDP_CHECK_CUDA(cudaMemcpyAsync(reinterpret_cast<void*>(dst[i]), reinterpret_cast<const void*>(src_host[i]),
                                   bytes[i], cudaMemcpyHostToDevice, s));
  }
  return 0;
}

}  // extern "C"
