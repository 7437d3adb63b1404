#include <stdlib.h>
// Host-side plumbing shared by all kernels: error string, device init, TMA descriptor encode.
#include "common.cuh"

#include <mutex>
#include <stdarg.h>
#include <string.h>

#include "raydp_b200.h"

namespace rdp {

static thread_local char g_err[1024] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("RAYDP_B200_PDL");
    on = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

static int g_num_sms[64] = {0};
int num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
  int n = __atomic_load_n(&g_num_sms[dev], __ATOMIC_RELAXED);
  if (n == 0) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    __atomic_store_n(&g_num_sms[dev], n, __ATOMIC_RELAXED);
  }
  return n;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

static void load_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  // resolved through the runtime so the library has no link-time dependency on libcuda
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
      q == cudaDriverEntryPointSuccess) {
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
}

int make_tmap_2d_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols,
                      uint64_t pitch_elems, uint32_t box_rows, uint32_t box_cols) {
  std::call_once(g_encode_once, load_encode);
  RDP_REQUIRE(g_encode != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  RDP_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p is not 16-byte aligned", base);
  RDP_REQUIRE((pitch_elems * 2) % 16 == 0, "TMA row pitch %llu elements is not a multiple of 16 bytes",
              (unsigned long long)pitch_elems);
  RDP_REQUIRE(box_cols * 2 <= 128 && box_rows <= 256, "TMA box %ux%u too large for 128B swizzle",
              box_rows, box_cols);
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch_elems * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RDP_REQUIRE(rc == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu "
              "pitch=%llu box=%ux%u)", int(rc), (unsigned long long)rows, (unsigned long long)cols,
              (unsigned long long)pitch_elems, box_rows, box_cols);
  return 0;
}

}  // namespace rdp

extern "C" {

const char* rdp_last_error() { return rdp::get_error(); }

int rdp_cuda_init(int device) {
  RDP_CHECK_CUDA(cudaSetDevice(device));
  RDP_CHECK_CUDA(cudaFree(0));
  int major = 0, minor = 0;
  RDP_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  RDP_CHECK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  RDP_REQUIRE(major == 10, "raydp_b200 kernels are built for sm_100a only; device %d is sm_%d%d",
              device, major, minor);
  return rdp::num_sms();
}

int rdp_num_sms() { return rdp::num_sms(); }

int rdp_memcpy_h2d_async(void* dst, const void* src_host, uint64_t bytes, void* stream) {
  RDP_CHECK_CUDA(cudaMemcpyAsync(dst, src_host, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
  return 0;
}

int rdp_memcpy_d2h_async(void* dst_host, const void* src, uint64_t bytes, void* stream) {
  RDP_CHECK_CUDA(cudaMemcpyAsync(dst_host, src, bytes, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
  return 0;
}

}  // extern "C"
