// Shared device/host helpers for the sm_100a kernels: error plumbing, PTX wrappers for
// mbarrier / TMA / tcgen05 / TMEM, UMMA descriptor builders, warp reductions.
//
// Descriptor encodings follow the PTX ISA "tcgen05 matrix descriptor" / "instruction
// descriptor" tables (cross-checked against cute/arch/mma_sm100_desc.hpp in the vendored
// CUTLASS headers); everything is written as inline PTX so the SASS shows UTCHMMA / UTMALDG /
// LDTM directly.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <string.h>
#include <stdint.h>
#include <stdio.h>

namespace rdp {

// ------------------------------------------------------------------------------------------
// host-side error state (one message per thread, surfaced through rdp_last_error())
// ------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
const char* get_error();

#define RDP_CHECK_CUDA(expr)                                                             \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      rdp::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,                \
                     cudaGetErrorString(_e));                                            \
      return -int(_e) - 1000;                                                            \
    }                                                                                    \
  } while (0)

#define RDP_REQUIRE(cond, ...)                                                           \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      rdp::set_error(__VA_ARGS__);                                                       \
      return -22;                                                                        \
    }                                                                                    \
  } while (0)

int num_sms();      // of the CURRENT device (cached per device)

// ---- programmatic dependent launch ------------------------------------------------------------------
// A training step is ~25 short kernels back to back inside one CUDA graph; each pays launch latency, a prologue
// (barrier init, TMEM allocation, descriptor prefetch) and a tail (last stores draining) -- ~7 us of a 25 us GEMM
// (profiles/gemm_tile_timeline.md).  Kernels launched with launch_pdl() may START while their predecessor in the
// stream is still finishing: everything before pdl_wait() overlaps the predecessor's tail, pdl_wait() returns once
// the predecessor has completed and its memory is visible.  Rules for a kernel launched this way:
//   * no global-memory access (loads, stores, TMA, atomics) before pdl_wait(), by any thread;
//   * pdl_trigger() early, so the NEXT kernel can do the same.
// Without the launch attribute both instructions are no-ops.  RAYDP_B200_PDL=0 launches everything normally.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();

template <typename Kernel, typename... Args>
inline cudaError_t launch_pdl(Kernel kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr;
  memset(&attr, 0, sizeof(attr));
  attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, args...);
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute of a kernel: a process that
// launches on a second GPU must set it there too.  `done` is one bit per device ordinal.
#define RDP_SMEM_ATTR_ONCE_PER_DEVICE(kernel, bytes)                                                    \
  do {                                                                                                  \
    static unsigned long long _done = 0ull;                                                             \
    int _dev = 0;                                                                                       \
    RDP_CHECK_CUDA(cudaGetDevice(&_dev));                                                               \
    if (_dev >= 64 || !((__atomic_load_n(&_done, __ATOMIC_ACQUIRE) >> _dev) & 1ull)) {                  \
      RDP_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)); \
      if (_dev < 64) __atomic_fetch_or(&_done, 1ull << _dev, __ATOMIC_RELEASE);                         \
    }                                                                                                   \
  } while (0)

// TMA descriptor for a row-major [rows, cols] bf16 matrix (pitch in elements), box
// [box_rows, box_cols], 128-byte swizzle.  Out-of-bounds elements read as zero.
int make_tmap_2d_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                      uint64_t pitch_elems, uint32_t box_rows, uint32_t box_cols);

// CTA-pair (cta_group::2) large-GEMM kernels, gemm_2cta.cu
int gemm2_tn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K,
             const float* bias, int relu, int mh, cudaStream_t stream);
int gemm2_nn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K, int mh,
             cudaStream_t stream);
int gemm2_wgrad(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw, int Nw,
                int mh, cudaStream_t stream);

#ifdef __CUDACC__
// ------------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// ---- mbarrier ----------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// ---- TMA ---------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
// 2-D tile load: coordinates are (col, row) = (innermost, outer).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar,
                                            int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)),
        "r"(c0), "r"(c1)
      : "memory");
}

// ---- tcgen05 / TMEM ----------------------------------------------------------------------
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {
  static_assert(kCols >= 32 && kCols <= 512 && (kCols & (kCols - 1)) == 0, "power of 2 in [32,512]");
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_result)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem], bf16 inputs, fp32 accumulate, issued by ONE thread.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once every previously issued MMA of this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets row (lane base + t).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---- UMMA descriptors ----------------------------------------------------------------------
// Shared-memory matrix descriptor, 128-byte swizzle.
//   K-major  operand tile: rows of 128 B (64 bf16 along K), 8-row groups 1024 B apart
//            -> SBO = 1024 B, LBO unused (1).
//   MN-major operand tile: rows of 128 B (64 bf16 along M/N), one row per k index, 8-k groups
//            1024 B apart -> SBO = 1024 B; successive 64-wide MN chunks LBO bytes apart.
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes,
                                                   uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= uint64_t((smem_addr & 0x3FFFF) >> 4);           // start address   [0,14)
  d |= uint64_t((lbo_bytes >> 4) & 0x3FFF) << 16;      // leading offset  [16,30)
  d |= uint64_t((sbo_bytes >> 4) & 0x3FFF) << 32;      // stride offset   [32,46)
  d |= uint64_t(1) << 46;                              // descriptor version (sm_100)
  d |= uint64_t(2) << 61;                              // SWIZZLE_128B
  return d;
}
// Instruction descriptor for kind::f16 with bf16 A/B and fp32 D.
__host__ __device__ __forceinline__ uint32_t umma_idesc_bf16(uint32_t m, uint32_t n, bool a_mn_major,
                                                             bool b_mn_major) {
  uint32_t d = 0;
  d |= 1u << 4;                      // D format  = f32
  d |= 1u << 7;                      // A format  = bf16
  d |= 1u << 10;                     // B format  = bf16
  d |= (a_mn_major ? 1u : 0u) << 15; // A major
  d |= (b_mn_major ? 1u : 0u) << 16; // B major
  d |= ((n >> 3) & 0x3F) << 17;      // N / 8
  d |= ((m >> 4) & 0x1F) << 24;      // M / 16
  return d;
}

// ---- misc ------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Column sums across the 32 lanes of a warp: on entry lane t holds v[0..31] = 32 columns of
// its row; on exit lane j holds sum over rows of column j.  31 shuffles, no shared memory.
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int half = 16; half >= 1; half >>= 1) {
    const bool upper = (lane & half) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      // keep the half of the columns this lane is responsible for, send the other half
      float keep = upper ? v[i + half] : v[i];
      float send = upper ? v[i] : v[i + half];
      float got = __shfl_xor_sync(0xffffffffu, send, half);
      v[i] = keep + got;
    }
  }
  return v[0];
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(h);
}
#endif  // __CUDACC__

}  // namespace rdp
