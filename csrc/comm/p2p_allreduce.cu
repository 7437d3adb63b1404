// K1 / K2: gradient all-reduce over NVLink peer memory, fused with the optimizer
// (SURVEY.md section 5.8, section 8 rows K1, K2, S2).
//
// Replaces the NCCL all-reduce that torch DDP issues for the reference
// (`train.torch.prepare_model` -> DDP, reference python/raydp/torch/estimator.py:243) plus the
// separate `optimizer.step()` (`:284-286`).  One process per GPU; every rank owns a *symmetric*
// cudaMalloc region that all peers map through CUDA IPC:
//
//     [ send[0] | send[1] | recv[0] | recv[1] | flags ]
//
//  * one-shot (small messages, e.g. the 48 513-parameter NYC-taxi MLP): copy the local gradient
//    into send[e&1], signal every peer, wait for every peer, then each rank pulls all W copies
//    straight over NVLink, sums in registers and -- in the fused variant -- applies 1/W scaling
//    and the Adam / SGD update in the same pass.  One kernel, one flag round.
//  * two-shot (large messages): rank r reduces slice r from all peers (reduce-scatter by pull)
//    and pushes the reduced slice into recv[e&1] of every peer (all-gather by push), a second
//    flag round publishes it, and the optimizer epilogue reads recv[] locally.
//
// Flags are monotonically increasing epochs (no reset races); buffers alternate on epoch parity
// so a rank can start filling the next step's buffer while a slow peer still reads this one.
// The epoch lives in device memory and is advanced by the kernel, so a captured CUDA graph can
// replay the step.  Blocks of one launch wait on other GPUs, never on each other's *launches*:
// the grid is capped so that all its blocks are resident at once.
//
// Single-GPU emulation (tests): gridDim.y = world, blockIdx.y = rank, all "peer" pointers local,
// launched cooperatively so every rank's blocks are co-resident.
#include "common.cuh"
#include "raydp_b200.h"

#include <cuda_bf16.h>
#include <string.h>

namespace rdp {

constexpr int kMaxRanks = 8;
constexpr int kArThreads = 512;

struct OptSpec {
  int kind;            // 0 none (write the sum * scale to out), 1 Adam/AdamW, 2 SGD
  float* p;            // parameters (fp32)
  float* m;            // Adam m / SGD momentum buffer
  float* v;            // Adam v
  float lr, beta1, beta2, eps, weight_decay, momentum, dampening;
  int adamw, nesterov;
  const float* lr_ptr;     // optional device scalars
  const float* step_ptr;
  __nv_bfloat16* shadow;   // optional bf16 copy of the updated parameters (GEMM operand), n elements
};

struct RankArgs {
  const float* in;                 // local gradient (n floats)
  float* out;                      // local result (may alias `in`)
  float* send[2][kMaxRanks];       // send[b][q]: rank q's send buffer, as mapped in THIS rank
  float* recv[2][kMaxRanks];       // recv[b][q]
  uint32_t* flags[kMaxRanks];      // flags[q]: rank q's flag array (2 * kMaxRanks words)
  uint32_t* epoch;                 // local: last completed epoch
  uint32_t* counters;              // local: [0] copy-done, [1] finished, [2] reduce-done
  OptSpec opt;
};

struct ArParams {
  RankArgs rank[kMaxRanks];        // indexed by blockIdx.y in emulation, [0] otherwise
  int rank_base;                   // this process' rank (0 in emulation)
  int world;
  int64_t n;                       // floats
  float scale;                     // applied to the sum (1/world for averaging)
  int fixed_buffer;                // two-shot only: always use send[0] / recv[0] (see rdp_allreduce, algo 3)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer(const float* p) {
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// ---- bounded waits -------------------------------------------------------------------------------
// A wait on a peer's flag must not outlive the peer: a rank that died (or never reached the collective)
// would otherwise leave this GPU spinning until the process is killed, and nothing above -- FailureConfig
// restarts, the trainer's failure detection -- could ever run.  Every spin checks %globaltimer now and
// then; past the deadline it records a code in a host-mapped status word, raises a device-wide "failed"
// flag (so every later wait of this process returns at once) and RETURNS: the kernel finishes with a
// meaningless result, the host sees the status at its next check and raises (raydp_b200.comm.check_status).
__device__ unsigned long long g_spin_timeout_ns = 0ull;     // 0 = wait forever
__device__ uint32_t* g_comm_status = nullptr;               // page-locked host word, device-mapped
__device__ int g_comm_failed = 0;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __noinline__ void report_wait_timeout(uint32_t code) {
  atomicExch(&g_comm_failed, 1);
  uint32_t* st = g_comm_status;
  if (st != nullptr) {
    *reinterpret_cast<volatile uint32_t*>(st) = code;
    __threadfence_system();
  }
}

// true when *f reached epoch e; false on timeout / after an earlier failure
__device__ __forceinline__ bool wait_epoch(const uint32_t* f, uint32_t e, uint32_t code) {
  uint32_t spins = 0;
  unsigned long long t0 = 0ull;
  while (int32_t(ld_acquire_sys(f) - e) < 0) {
    if ((++spins & 1023u) == 0u) {
      if (*reinterpret_cast<volatile int*>(&g_comm_failed)) return false;
      const unsigned long long lim = g_spin_timeout_ns;
      if (lim) {
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0ull) t0 = now;
        else if (now - t0 > lim) { report_wait_timeout(code); return false; }
      }
    }
  }
  return true;
}

// All blocks of this rank's grid have finished their part -> the last one signals `slot` on every
// peer; then every block waits until all ranks have signalled epoch e on OUR flag array.
__device__ __forceinline__ void grid_signal_and_wait(const RankArgs& a, int my_rank, int world, uint32_t e,
                                                     int slot, uint32_t* counter, int nblocks) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(counter, 1u);
    if (prev == uint32_t(nblocks - 1)) {
      *counter = 0;
      __threadfence_system();
      for (int q = 0; q < world; ++q) st_release_sys(a.flags[q] + slot * kMaxRanks + my_rank, e);
    }
  }
  if (int(threadIdx.x) < world) {
    const uint32_t* f = a.flags[my_rank] + slot * kMaxRanks + threadIdx.x;
    wait_epoch(f, e, 0x100u + uint32_t(slot));
  }
  __syncthreads();
}

__device__ __forceinline__ void apply_update(const OptSpec& o, int64_t i, float4 g, float bc1, float bc2, float lr,
                                             bool first_step, float* out) {
  float gs[4] = {g.x, g.y, g.z, g.w};
  if (o.kind == 0) {
    *reinterpret_cast<float4*>(out + i) = g;
    return;
  }
  float4 p4 = *reinterpret_cast<float4*>(o.p + i);
  float ps[4] = {p4.x, p4.y, p4.z, p4.w};
  if (o.kind == 1) {
    float4 m4 = *reinterpret_cast<float4*>(o.m + i), v4 = *reinterpret_cast<float4*>(o.v + i);
    float ms[4] = {m4.x, m4.y, m4.z, m4.w}, vs[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gk = gs[k];
      if (o.weight_decay != 0.f) { if (o.adamw) ps[k] *= (1.f - lr * o.weight_decay); else gk += o.weight_decay * ps[k]; }
      ms[k] = o.beta1 * ms[k] + (1.f - o.beta1) * gk;
      vs[k] = o.beta2 * vs[k] + (1.f - o.beta2) * gk * gk;
      ps[k] -= (lr / bc1) * (ms[k] / (sqrtf(vs[k]) / sqrtf(bc2) + o.eps));
    }
    *reinterpret_cast<float4*>(o.m + i) = make_float4(ms[0], ms[1], ms[2], ms[3]);
    *reinterpret_cast<float4*>(o.v + i) = make_float4(vs[0], vs[1], vs[2], vs[3]);
  } else {
    float4 b4 = o.m ? *reinterpret_cast<float4*>(o.m + i) : make_float4(0, 0, 0, 0);
    float bs[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float gk = gs[k] + o.weight_decay * ps[k];
      if (o.m != nullptr && o.momentum != 0.f) {
        bs[k] = first_step ? gk : o.momentum * bs[k] + (1.f - o.dampening) * gk;
        gk = o.nesterov ? gk + o.momentum * bs[k] : bs[k];
      }
      ps[k] -= lr * gk;
    }
    if (o.m) *reinterpret_cast<float4*>(o.m + i) = make_float4(bs[0], bs[1], bs[2], bs[3]);
  }
  *reinterpret_cast<float4*>(o.p + i) = make_float4(ps[0], ps[1], ps[2], ps[3]);
  if (o.shadow != nullptr) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(ps[0], ps[1]), hi = __floats2bfloat162_rn(ps[2], ps[3]);
    *reinterpret_cast<uint2*>(o.shadow + i) = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  }
  if (out != nullptr) *reinterpret_cast<float4*>(out + i) = g;
}

template <bool kTwoShot>
__global__ void __launch_bounds__(kArThreads)
allreduce_kernel(const __grid_constant__ ArParams P) {
  const RankArgs& a = P.rank[blockIdx.y];
  const int my_rank = P.rank_base + blockIdx.y;
  const int world = P.world;
  const int nblocks = gridDim.x;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(a.epoch) + 1u;
  const int b = P.fixed_buffer ? 0 : int(e & 1);
  const int64_t n4 = P.n / 4;                      // n is padded to a multiple of 4 by the host
  const int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  const int64_t nthreads = int64_t(nblocks) * blockDim.x;

  // optimizer scalars (device-resident so CUDA graphs follow lr schedules / bias correction)
  float lr = a.opt.lr, bc1 = 1.f, bc2 = 1.f;
  bool first_step = false;
  if (a.opt.kind) {
    if (a.opt.lr_ptr) lr = *a.opt.lr_ptr;
    const float t = a.opt.step_ptr ? *a.opt.step_ptr : 1.f;
    bc1 = 1.f - powf(a.opt.beta1, t);
    bc2 = 1.f - powf(a.opt.beta2, t);
    first_step = t <= 1.f;
  }

  // phase 1: publish my gradient in my send buffer
  float* mine = a.send[b][my_rank];
  if (a.in != mine) {      // registered mode: the gradient was produced straight into the send buffer
    for (int64_t i = tid; i < n4; i += nthreads)
      reinterpret_cast<float4*>(mine)[i] = reinterpret_cast<const float4*>(a.in)[i];
  }
  grid_signal_and_wait(a, my_rank, world, e, 0, a.counters + 0, nblocks);

  if (!kTwoShot) {
    // phase 2: pull every rank's copy over NVLink, sum in registers, update.  The sum runs in rank
    // order on every rank so that all replicas compute bit-identical parameters.
    for (int64_t i = tid; i < n4; i += nthreads) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
      for (int q = 0; q < world; ++q) {
        const float4 v = ld_peer(a.send[b][q] + i * 4);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      s.x *= P.scale; s.y *= P.scale; s.z *= P.scale; s.w *= P.scale;
      apply_update(a.opt, i * 4, s, bc1, bc2, lr, first_step, a.out);
    }
  } else {
    // phase 2: reduce-scatter by pull (slice my_rank), all-gather by push into every recv[b]
    const int64_t per = (n4 + world - 1) / world;
    const int64_t lo = per * my_rank, hi = min(n4, lo + per);
    for (int64_t i0 = lo + tid; i0 < hi; i0 += 2 * nthreads) {
      const int64_t i1 = i0 + nthreads;
      const bool two = i1 < hi;
      float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
#pragma unroll 4
      for (int q = 0; q < world; ++q) {
        const float* src = a.send[b][(my_rank + q) % world];
        const float4 v0 = ld_peer(src + i0 * 4);
        float4 v1 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (two) v1 = ld_peer(src + i1 * 4);
        s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
        s1.x += v1.x; s1.y += v1.y; s1.z += v1.z; s1.w += v1.w;
      }
      s0.x *= P.scale; s0.y *= P.scale; s0.z *= P.scale; s0.w *= P.scale;
      s1.x *= P.scale; s1.y *= P.scale; s1.z *= P.scale; s1.w *= P.scale;
#pragma unroll 4
      for (int q = 0; q < world; ++q) {
        float4* dst = reinterpret_cast<float4*>(a.recv[b][(my_rank + q) % world]);
        dst[i0] = s0;
        if (two) dst[i1] = s1;
      }
    }
    grid_signal_and_wait(a, my_rank, world, e, 1, a.counters + 2, nblocks);
    // phase 3: optimizer epilogue (or copy-out) from the locally complete recv buffer
    const float* full = a.recv[b][my_rank];
    if (a.opt.kind != 0 || (a.out != nullptr && a.out != full)) {   // else: the caller reads recv[] in place
      for (int64_t i = tid; i < n4; i += nthreads)
        apply_update(a.opt, i * 4, reinterpret_cast<const float4*>(full)[i], bc1, bc2, lr, first_step, a.out);
    }
  }

  // the last block to finish advances the epoch for the next launch
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(a.counters + 1, 1u);
    if (prev == uint32_t(nblocks - 1)) {
      a.counters[1] = 0;
      *a.epoch = e;
    }
  }
}

// ---- bf16 on the wire, fp32 accumulate (large gradient buckets) -------------------------------------
// Two-shot all-reduce whose NVLink traffic is bf16: the fp32 gradient is rounded once into the bf16 send
// buffer, every rank pulls its slice of all W copies (16 bytes = 8 values per load), sums them in fp32 in a
// fixed rank order, scales, rounds the average once and pushes it as bf16 into every rank's receive buffer;
// the optimizer epilogue reads the local bf16 copy.  Half the link bytes of the fp32 kernel; the same
// precision contract as DDP's bf16 compression hook, with the sum kept in fp32 (python/raydp/torch/estimator.py
// trains through DDP's fp32 buckets -- this is an opt-in trade for link-bound buckets).
__device__ __forceinline__ uint4 ld_peer_b128(const void* p) {
  uint4 v;
  asm volatile("ld.global.relaxed.sys.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void acc_bf16x8(float (&s)[8], const uint4 v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    s[2 * k] += __uint_as_float(w[k] << 16);
    s[2 * k + 1] += __uint_as_float(w[k] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

__global__ void __launch_bounds__(kArThreads)
allreduce_bf16_kernel(const __grid_constant__ ArParams P) {
  const RankArgs& a = P.rank[blockIdx.y];
  const int my_rank = P.rank_base + blockIdx.y;
  const int world = P.world;
  const int nblocks = gridDim.x;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(a.epoch) + 1u;
  const int64_t n8 = P.n / 8;                      // n is a multiple of 8 (host)
  const int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  const int64_t nthreads = int64_t(nblocks) * blockDim.x;
  float lr = a.opt.lr, bc1 = 1.f, bc2 = 1.f;
  bool first_step = false;
  if (a.opt.kind) {
    if (a.opt.lr_ptr) lr = *a.opt.lr_ptr;
    const float t = a.opt.step_ptr ? *a.opt.step_ptr : 1.f;
    bc1 = 1.f - powf(a.opt.beta1, t);
    bc2 = 1.f - powf(a.opt.beta2, t);
    first_step = t <= 1.f;
  }
  // phase 1: round my gradient into my bf16 send buffer
  uint4* mine = reinterpret_cast<uint4*>(a.send[0][my_rank]);
  for (int64_t i = tid; i < n8; i += nthreads) {
    const float4 g0 = reinterpret_cast<const float4*>(a.in)[2 * i], g1 = reinterpret_cast<const float4*>(a.in)[2 * i + 1];
    mine[i] = make_uint4(pack2_bf16(g0.x, g0.y), pack2_bf16(g0.z, g0.w), pack2_bf16(g1.x, g1.y), pack2_bf16(g1.z, g1.w));
  }
  grid_signal_and_wait(a, my_rank, world, e, 0, a.counters + 0, nblocks);
  // phase 2: reduce-scatter by pull (slice my_rank, fp32 sums in rank order), all-gather by bf16 push
  const int64_t per = (n8 + world - 1) / world;
  const int64_t lo = per * my_rank, hi = min(n8, lo + per);
  for (int64_t i = lo + tid; i < hi; i += nthreads) {
    float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    uint4 v[kMaxRanks];
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < world) v[q] = ld_peer_b128(reinterpret_cast<const uint4*>(a.send[0][q]) + i);
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < world) acc_bf16x8(s, v[q]);
    const uint4 r = make_uint4(pack2_bf16(s[0] * P.scale, s[1] * P.scale), pack2_bf16(s[2] * P.scale, s[3] * P.scale),
                               pack2_bf16(s[4] * P.scale, s[5] * P.scale), pack2_bf16(s[6] * P.scale, s[7] * P.scale));
#pragma unroll 4
    for (int q = 0; q < world; ++q) reinterpret_cast<uint4*>(a.recv[0][(my_rank + q) % world])[i] = r;
  }
  grid_signal_and_wait(a, my_rank, world, e, 1, a.counters + 2, nblocks);
  // phase 3: optimizer epilogue (or fp32 copy-out) from the locally complete bf16 receive buffer
  const uint4* full = reinterpret_cast<const uint4*>(a.recv[0][my_rank]);
  for (int64_t i = tid; i < n8; i += nthreads) {
    float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    acc_bf16x8(s, full[i]);
    apply_update(a.opt, i * 8, make_float4(s[0], s[1], s[2], s[3]), bc1, bc2, lr, first_step, a.out);
    apply_update(a.opt, i * 8 + 4, make_float4(s[4], s[5], s[6], s[7]), bc1, bc2, lr, first_step, a.out);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(a.counters + 1, 1u);
    if (prev == uint32_t(nblocks - 1)) {
      a.counters[1] = 0;
      *a.epoch = e;
    }
  }
}

// ---- LL: one-round all-reduce for latency-bound messages (<= ~1 MB) ---------------------------------
// The one-shot kernel above needs a flag round (copy -> signal -> wait) before anybody may pull.  Here the
// flag travels WITH the data: rank r pushes its gradient into slot r of every peer's receive area as 16-byte
// packets {v0, e, v1, e} (two floats, each sharing an 8-byte word with the epoch number e, so a packet that
// arrives torn in halves is still self-validating), then polls ITS OWN receive area until the packets of all
// ranks carry e, sums them in rank order (bit-identical replicas) and applies the optimizer.  No signal
// round, no inter-block synchronisation, twice the bytes -- the right trade below ~1 MB on NVSwitch.
// Receive areas alternate on epoch parity: a rank can be at most one collective ahead of the slowest rank
// (it needs that rank's packets to finish), so the area of epoch e is never overwritten before every rank
// has consumed it.  Layout of a rank's region (floats): ll[parity 0..1][source rank][2 * slot_floats].
struct LlRankArgs {
  const float* in;
  float* out;
  float* base[kMaxRanks];          // base[q]: rank q's region as mapped here
  uint32_t* epoch;
  uint32_t* counters;              // [1] finished
  OptSpec opt;
};

struct LlParams {
  LlRankArgs rank[kMaxRanks];
  int rank_base, world;
  int64_t n;                       // floats, multiple of 4
  int64_t slot_floats;
  float scale;
};

__device__ __forceinline__ void st_packet(float* p, float v0, float v1, uint32_t e) {
  asm volatile("st.relaxed.sys.global.v4.b32 [%0], {%1, %2, %3, %2};"
               ::"l"(p), "r"(__float_as_uint(v0)), "r"(e), "r"(__float_as_uint(v1)) : "memory");
}
__device__ __forceinline__ uint4 ld_packet(const float* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kArThreads)
allreduce_ll_kernel(const __grid_constant__ LlParams P) {
  const LlRankArgs& a = P.rank[blockIdx.y];
  const int my_rank = P.rank_base + blockIdx.y;
  const int world = P.world;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(a.epoch) + 1u;
  const int64_t area = 2 * P.slot_floats;                       // floats per (parity, source) slot
  const int64_t par_off = int64_t(e & 1u) * kMaxRanks * area;   // slots are laid out for kMaxRanks sources
  const int64_t n4 = P.n / 4;
  const int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  const int64_t nthreads = int64_t(gridDim.x) * blockDim.x;
  float lr = a.opt.lr, bc1 = 1.f, bc2 = 1.f;
  bool first_step = false;
  if (a.opt.kind) {
    if (a.opt.lr_ptr) lr = *a.opt.lr_ptr;
    const float t = a.opt.step_ptr ? *a.opt.step_ptr : 1.f;
    bc1 = 1.f - powf(a.opt.beta1, t);
    bc2 = 1.f - powf(a.opt.beta2, t);
    first_step = t <= 1.f;
  }
  // push: my four floats -> two packets in slot `my_rank` of every rank (own region included)
  for (int64_t i = tid; i < n4; i += nthreads) {
    const float4 g = reinterpret_cast<const float4*>(a.in)[i];
#pragma unroll 8
    for (int q = 0; q < world; ++q) {
      float* dst = a.base[(my_rank + q) % world] + par_off + int64_t(my_rank) * area + i * 8;
      st_packet(dst, g.x, g.y, e);
      st_packet(dst + 4, g.z, g.w, e);
    }
  }
  // pull locally: wait for every rank's packets, sum in rank order
  bool ok = true;
  for (int64_t i = tid; i < n4; i += nthreads) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int q = 0; q < world; ++q) {
      const float* src = a.base[my_rank] + par_off + int64_t(q) * area + i * 8;
      uint4 p0 = ld_packet(src), p1 = ld_packet(src + 4);
      uint32_t spins = 0;
      unsigned long long t0 = 0ull;
      while (ok && (p0.y != e || p0.w != e || p1.y != e || p1.w != e)) {
        p0 = ld_packet(src);
        p1 = ld_packet(src + 4);
        if ((++spins & 1023u) == 0u) {
          if (*reinterpret_cast<volatile int*>(&g_comm_failed)) { ok = false; break; }
          const unsigned long long lim = g_spin_timeout_ns;
          if (lim) {
            const unsigned long long now = globaltimer_ns();
            if (t0 == 0ull) t0 = now;
            else if (now - t0 > lim) { report_wait_timeout(0x400u + uint32_t(q)); ok = false; }
          }
        }
      }
      s.x += __uint_as_float(p0.x); s.y += __uint_as_float(p0.z);
      s.z += __uint_as_float(p1.x); s.w += __uint_as_float(p1.z);
    }
    s.x *= P.scale; s.y *= P.scale; s.z *= P.scale; s.w *= P.scale;
    apply_update(a.opt, i * 4, s, bc1, bc2, lr, first_step, a.out);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(a.counters + 1, 1u);
    if (prev == uint32_t(gridDim.x - 1)) {
      a.counters[1] = 0;
      *a.epoch = e;
    }
  }
}

// One flag round between all ranks (slot 2 of the flag area): everything this GPU wrote before the
// barrier kernel started -- including peer stores of earlier kernels on the stream -- is ordered
// before the flag, and the kernel returns only after every rank has signalled the same epoch.
struct BarrierParams {
  uint32_t* flags[kMaxRanks];
  int rank, world;
  uint32_t* epoch;
};

__global__ void barrier_kernel(const __grid_constant__ BarrierParams p) {
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(p.epoch) + 1u;
  __threadfence_system();
  if (int(threadIdx.x) < p.world) {
    st_release_sys(p.flags[threadIdx.x] + 2 * kMaxRanks + p.rank, e);
    const uint32_t* f = p.flags[p.rank] + 2 * kMaxRanks + threadIdx.x;
    wait_epoch(f, e, 0x200u);
  }
  __syncthreads();
  if (threadIdx.x == 0) *p.epoch = e;
}

}  // namespace rdp

using namespace rdp;

extern "C" {

int rdp_comm_barrier(void* const* bases, uint64_t flag_offset_bytes, int rank, int world, uint32_t* epoch,
                     void* stream) {
  RDP_REQUIRE(world >= 1 && world <= kMaxRanks, "rdp_comm_barrier: world=%d", world);
  BarrierParams p;
  for (int q = 0; q < world; ++q)
    p.flags[q] = reinterpret_cast<uint32_t*>(static_cast<char*>(bases[q]) + flag_offset_bytes);
  p.rank = rank; p.world = world; p.epoch = epoch;
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- symmetric memory plumbing (host) -----------------------------------------------------------
int rdp_comm_alloc(uint64_t bytes, void** out) {
  RDP_CHECK_CUDA(cudaMalloc(out, bytes));
  RDP_CHECK_CUDA(cudaMemset(*out, 0, bytes));
  RDP_CHECK_CUDA(cudaDeviceSynchronize());
  return 0;
}

int rdp_comm_free(void* ptr) {
  RDP_CHECK_CUDA(cudaFree(ptr));
  return 0;
}

int rdp_comm_ipc_handle(void* ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  RDP_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}

int rdp_comm_ipc_open(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  RDP_CHECK_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int rdp_comm_ipc_close(void* ptr) {
  RDP_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int rdp_comm_can_access_peer(int dev, int peer) {
  int ok = 0;
  if (cudaDeviceCanAccessPeer(&ok, dev, peer) != cudaSuccess) return 0;
  return ok;
}

// ---- all-reduce -----------------------------------------------------------------------------------
// Layout of one rank's symmetric region, in floats: send[0] | send[1] | recv[0] | recv[1] | flags.
// `bases[q]` = rank q's region as mapped in this process.  In emulation (n_local_ranks == world)
// every per-rank array has `world` entries and all pointers are local.
int rdp_allreduce(void* const* bases, uint64_t slot_floats, int rank, int world, int n_local_ranks,
                  const float* const* ins, float* const* outs, uint32_t* const* epochs,
                  uint32_t* const* counters, const rdp_opt_args* opts, int64_t n, float scale,
                  int algo, void* stream) {
  RDP_REQUIRE(world >= 1 && world <= kMaxRanks, "rdp_allreduce: world=%d (max %d)", world, kMaxRanks);
  RDP_REQUIRE(n_local_ranks == 1 || n_local_ranks == world, "rdp_allreduce: bad n_local_ranks");
  RDP_REQUIRE(n % 4 == 0, "rdp_allreduce: n=%lld must be a multiple of 4 (pad the flat buffer)", (long long)n);
  RDP_REQUIRE(n > 0 && uint64_t(n) <= slot_floats && slot_floats % 4 == 0,
              "rdp_allreduce: n=%lld exceeds the symmetric slot (%llu floats)", (long long)n,
              (unsigned long long)slot_floats);
  ArParams P;
  memset(&P, 0, sizeof(P));
  P.rank_base = n_local_ranks == 1 ? rank : 0;
  P.world = world;
  P.n = (n + 3) / 4 * 4;
  P.scale = scale;
  for (int l = 0; l < n_local_ranks; ++l) {
    RankArgs& a = P.rank[l];
    a.in = ins[l];
    a.out = outs[l];
    a.epoch = epochs[l];
    a.counters = counters[l];
    for (int q = 0; q < world; ++q) {
      float* base = static_cast<float*>(bases[q]);
      a.send[0][q] = base;
      a.send[1][q] = base + slot_floats;
      a.recv[0][q] = base + 2 * slot_floats;
      a.recv[1][q] = base + 3 * slot_floats;
      a.flags[q] = reinterpret_cast<uint32_t*>(base + 4 * slot_floats);
    }
    if (opts != nullptr) {
      const rdp_opt_args& o = opts[l];
      a.opt = OptSpec{o.kind, o.p, o.m, o.v, o.lr, o.beta1, o.beta2, o.eps, o.weight_decay, o.momentum,
                      o.dampening, o.adamw, o.nesterov, o.lr_ptr, o.step_ptr,
                      static_cast<__nv_bfloat16*>(o.shadow)};
    }
  }
  // algo 3: two-shot on send[0]/recv[0] only.  Safe without parity: the second flag round ends after every
  // rank has finished reading all send buffers, and recv is rewritten only after the next call's first round.
  const bool two_shot = algo == 2 || algo == 3 || (algo == 0 && n * 4 > (1 << 20));
  P.fixed_buffer = algo == 3 ? 1 : 0;
  int64_t work = two_shot ? P.n / 4 / world : P.n / 4;
  int blocks = int((work + kArThreads - 1) / kArThreads);
  int cap = two_shot ? num_sms() : 32;
  if (n_local_ranks > 1) cap = max(1, (num_sms() / 2) / world);      // emulation: all ranks share one GPU
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  dim3 grid(blocks, n_local_ranks);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  void* args[] = {&P};
  if (n_local_ranks > 1) {
    // ranks of an emulated world wait on each other inside one grid: co-residency must be guaranteed
    if (two_shot) RDP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)allreduce_kernel<true>, grid, dim3(kArThreads), args, 0, s));
    else RDP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)allreduce_kernel<false>, grid, dim3(kArThreads), args, 0, s));
  } else {
    if (two_shot) allreduce_kernel<true><<<grid, kArThreads, 0, s>>>(P);
    else allreduce_kernel<false><<<grid, kArThreads, 0, s>>>(P);
  }
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// bf16-wire two-shot all-reduce (algo 5): same region layout and arguments as rdp_allreduce; the send / recv
// buffers are used as bf16 (send[0] / recv[0] only, as algo 3).  n % 8 == 0; `ins` must not alias the region.
int rdp_allreduce_bf16(void* const* bases, uint64_t slot_floats, int rank, int world, int n_local_ranks,
                       const float* const* ins, float* const* outs, uint32_t* const* epochs,
                       uint32_t* const* counters, const rdp_opt_args* opts, int64_t n, float scale, void* stream) {
  RDP_REQUIRE(world >= 1 && world <= kMaxRanks, "rdp_allreduce_bf16: world=%d (max %d)", world, kMaxRanks);
  RDP_REQUIRE(n_local_ranks == 1 || n_local_ranks == world, "rdp_allreduce_bf16: bad n_local_ranks");
  RDP_REQUIRE(n > 0 && n % 8 == 0 && uint64_t(n) <= 2 * slot_floats && slot_floats % 4 == 0,
              "rdp_allreduce_bf16: n=%lld must be a multiple of 8 within the slot (%llu bf16 values)", (long long)n,
              (unsigned long long)(2 * slot_floats));
  ArParams P;
  memset(&P, 0, sizeof(P));
  P.rank_base = n_local_ranks == 1 ? rank : 0;
  P.world = world; P.n = n; P.scale = scale; P.fixed_buffer = 1;
  for (int l = 0; l < n_local_ranks; ++l) {
    RankArgs& a = P.rank[l];
    a.in = ins[l]; a.out = outs[l]; a.epoch = epochs[l]; a.counters = counters[l];
    for (int q = 0; q < world; ++q) {
      float* base = static_cast<float*>(bases[q]);
      a.send[0][q] = base;
      a.send[1][q] = base + slot_floats;
      a.recv[0][q] = base + 2 * slot_floats;
      a.recv[1][q] = base + 3 * slot_floats;
      a.flags[q] = reinterpret_cast<uint32_t*>(base + 4 * slot_floats);
      RDP_REQUIRE(q != P.rank_base + l || static_cast<const void*>(a.in) != static_cast<const void*>(base),
                  "rdp_allreduce_bf16: the gradient may not live in the send buffer (it is rewritten as bf16)");
    }
    if (opts != nullptr) {
      const rdp_opt_args& o = opts[l];
      a.opt = OptSpec{o.kind, o.p, o.m, o.v, o.lr, o.beta1, o.beta2, o.eps, o.weight_decay, o.momentum,
                      o.dampening, o.adamw, o.nesterov, o.lr_ptr, o.step_ptr,
                      static_cast<__nv_bfloat16*>(o.shadow)};
    }
  }
  int64_t work = n / 8 / world;
  int blocks = int((work + kArThreads - 1) / kArThreads);
  int cap = num_sms();
  if (n_local_ranks > 1) cap = max(1, (num_sms() / 2) / world);
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  dim3 grid(blocks, n_local_ranks);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  void* args[] = {&P};
  if (n_local_ranks > 1) {
    RDP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)allreduce_bf16_kernel, grid, dim3(kArThreads), args, 0, s));
  } else {
    allreduce_bf16_kernel<<<grid, kArThreads, 0, s>>>(P);
  }
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// LL one-round all-reduce (latency-bound sizes).  Region of one rank: 2 * 8 * 2 * slot_floats floats
// (parity x source-rank slots x {value, epoch} packets); zero-initialised; epochs[l] / counters[l] as for
// rdp_allreduce.  n <= slot_floats, n % 4 == 0.
int rdp_allreduce_ll(void* const* bases, uint64_t slot_floats, int rank, int world, int n_local_ranks,
                     const float* const* ins, float* const* outs, uint32_t* const* epochs,
                     uint32_t* const* counters, const rdp_opt_args* opts, int64_t n, float scale, void* stream) {
  RDP_REQUIRE(world >= 1 && world <= kMaxRanks, "rdp_allreduce_ll: world=%d (max %d)", world, kMaxRanks);
  RDP_REQUIRE(n_local_ranks == 1 || n_local_ranks == world, "rdp_allreduce_ll: bad n_local_ranks");
  RDP_REQUIRE(n > 0 && n % 4 == 0 && uint64_t(n) <= slot_floats && slot_floats % 4 == 0,
              "rdp_allreduce_ll: n=%lld must be a multiple of 4 within the slot (%llu floats)", (long long)n,
              (unsigned long long)slot_floats);
  LlParams P;
  memset(&P, 0, sizeof(P));
  P.rank_base = n_local_ranks == 1 ? rank : 0;
  P.world = world; P.n = n; P.slot_floats = int64_t(slot_floats); P.scale = scale;
  for (int l = 0; l < n_local_ranks; ++l) {
    LlRankArgs& a = P.rank[l];
    a.in = ins[l]; a.out = outs[l]; a.epoch = epochs[l]; a.counters = counters[l];
    for (int q = 0; q < world; ++q) a.base[q] = static_cast<float*>(bases[q]);
    if (opts != nullptr) {
      const rdp_opt_args& o = opts[l];
      a.opt = OptSpec{o.kind, o.p, o.m, o.v, o.lr, o.beta1, o.beta2, o.eps, o.weight_decay, o.momentum,
                      o.dampening, o.adamw, o.nesterov, o.lr_ptr, o.step_ptr,
                      static_cast<__nv_bfloat16*>(o.shadow)};
    }
  }
  int blocks = int((n / 4 + kArThreads - 1) / kArThreads);
  int cap = num_sms();
  if (n_local_ranks > 1) cap = max(1, (num_sms() / 2) / world);
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  dim3 grid(blocks, n_local_ranks);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  void* args[] = {&P};
  if (n_local_ranks > 1) {
    RDP_CHECK_CUDA(cudaLaunchCooperativeKernel((void*)allreduce_ll_kernel, grid, dim3(kArThreads), args, 0, s));
  } else {
    allreduce_ll_kernel<<<grid, kArThreads, 0, s>>>(P);
  }
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Bounded peer waits: after `seconds` (0 = forever) a wait on a peer's flag gives up, stores a code (!= 0)
// into *host_status (page-locked host memory) and every later wait of this process returns at once.
int rdp_comm_set_timeout(double seconds, uint32_t* host_status) {
  unsigned long long ns = seconds > 0 ? (unsigned long long)(seconds * 1e9) : 0ull;
  void* mapped = nullptr;
  if (host_status != nullptr) RDP_CHECK_CUDA(cudaHostGetDevicePointer(&mapped, host_status, 0));
  uint32_t* st = static_cast<uint32_t*>(mapped);
  int zero = 0;
  RDP_CHECK_CUDA(cudaMemcpyToSymbol(g_spin_timeout_ns, &ns, sizeof(ns)));
  RDP_CHECK_CUDA(cudaMemcpyToSymbol(g_comm_status, &st, sizeof(st)));
  RDP_CHECK_CUDA(cudaMemcpyToSymbol(g_comm_failed, &zero, sizeof(zero)));
  return 0;
}

}  // extern "C"

// ================================================================================================
// K3: NVLS all-reduce -- the reduction happens inside the NVSwitch
// ================================================================================================
// `mc` is a *multicast* mapping of every rank's symmetric buffer (created with
// torch.distributed._symmetric_memory, which only does the cuMulticast* / handle-exchange
// plumbing; the kernel below is ours).  Rank r owns slice r: one `multimem.ld_reduce` pulls the
// sum of all W copies of a 16-byte vector through the switch, one `multimem.st` broadcasts the
// result into all W buffers.  Per rank that is S/W bytes in and S/W bytes out instead of the
// 2 (W-1)/W * S of the pull/push two-shot, so large messages are no longer bound by this GPU's
// link bandwidth.  Two flag rounds delimit the kernel; the result is in place in every buffer.
namespace rdp {

struct NvlsParams {
  float* mc;                       // multicast address of the buffer
  float* local;                    // this rank's unicast address (optimizer epilogue reads it)
  uint32_t* flags[kMaxRanks];      // peer-mapped flag words (unicast), 4 slots x kMaxRanks
  int rank, world;
  int64_t n;                       // floats, multiple of 4
  float scale;
  uint32_t* epoch;
  uint32_t* counters;              // [0] arrive-A, [1] finished, [2] arrive-B
  OptSpec opt;
  float* out;                      // optional separate output (null: leave the result in `local`)
};

__device__ __forceinline__ void flags_signal_and_wait(uint32_t* const* flags, int my_rank, int world, uint32_t e,
                                                      int slot, uint32_t* counter, int nblocks) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(counter, 1u);
    if (prev == uint32_t(nblocks - 1)) {
      *counter = 0;
      __threadfence_system();
      for (int q = 0; q < world; ++q) st_release_sys(flags[q] + slot * kMaxRanks + my_rank, e);
    }
  }
  if (int(threadIdx.x) < world) {
    const uint32_t* f = flags[my_rank] + slot * kMaxRanks + threadIdx.x;
    wait_epoch(f, e, 0x300u + uint32_t(slot));
  }
  __syncthreads();
}

__device__ __forceinline__ void nvls_allreduce_body(const NvlsParams& p) {
  const int nblocks = gridDim.x;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(p.epoch) + 1u;
  const int64_t n4 = p.n / 4;
  const int64_t per = (n4 + p.world - 1) / p.world;
  const int64_t lo = per * p.rank, hi = min(n4, lo + per);
  const int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  const int64_t nthreads = int64_t(nblocks) * blockDim.x;
  float lr = p.opt.lr, bc1 = 1.f, bc2 = 1.f;
  bool first_step = false;
  if (p.opt.kind) {
    if (p.opt.lr_ptr) lr = *p.opt.lr_ptr;
    const float t = p.opt.step_ptr ? *p.opt.step_ptr : 1.f;
    bc1 = 1.f - powf(p.opt.beta1, t);
    bc2 = 1.f - powf(p.opt.beta2, t);
    first_step = t <= 1.f;
  }
  // round A: every rank's contribution is in its buffer
  flags_signal_and_wait(p.flags, p.rank, p.world, e, 0, p.counters + 0, nblocks);
  // four independent switch reductions in flight per thread before the first broadcast store
  constexpr int kU = 4;
  for (int64_t i0 = lo + tid; i0 < hi; i0 += nthreads * kU) {
    float4 v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t i = i0 + u * nthreads;
      if (i < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(p.mc + i * 4) : "memory");
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t i = i0 + u * nthreads;
      if (i < hi)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(p.mc + i * 4), "f"(v[u].x * p.scale), "f"(v[u].y * p.scale), "f"(v[u].z * p.scale),
                       "f"(v[u].w * p.scale) : "memory");
    }
  }
  // round B: every slice has been broadcast into every buffer
  flags_signal_and_wait(p.flags, p.rank, p.world, e, 1, p.counters + 2, nblocks);
  if (p.opt.kind || p.out != nullptr) {
    for (int64_t i = tid; i < n4; i += nthreads)
      apply_update(p.opt, i * 4, reinterpret_cast<const float4*>(p.local)[i], bc1, bc2, lr, first_step, p.out);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(p.counters + 1, 1u);
    if (prev == uint32_t(nblocks - 1)) {
      p.counters[1] = 0;
      *p.epoch = e;
    }
  }
}

__global__ void __launch_bounds__(kArThreads)
nvls_allreduce_kernel(const __grid_constant__ NvlsParams p) { nvls_allreduce_body(p); }

// Overlap shape: 256 threads and at most 40 registers per thread, one block per SM.  That is
// 10 K registers, which fits NEXT TO a resident GEMM CTA (the 512-row pair kernel holds 320 x 168 =
// 54 K of the SM's 64 K registers), so the all-reduce + optimizer of one gradient bucket really
// runs concurrently with the backward GEMMs of the next layers instead of queueing behind their
// persistent CTAs.  148 x 256 threads x 4 x 16 B in flight is still enough to fill the NVLink pipe.
constexpr int kArOverlapThreads = 256;
// (min 6 blocks/SM in the launch bounds is how the 40-register cap is expressed)
__global__ void __launch_bounds__(kArOverlapThreads, 6)
nvls_allreduce_overlap_kernel(const __grid_constant__ NvlsParams p) { nvls_allreduce_body(p); }

// bf16 on the wire through the switch: the multicast buffer holds bf16.  The fp32 gradient `in` is rounded into
// the local buffer, `multimem.ld_reduce ... .acc::f32.v4.bf16x2` sums the W copies of 8 values in the switch with
// fp32 accumulation, the scaled average is rounded once and broadcast with `multimem.st`; the optimizer epilogue
// reads the local bf16 copy.  Half the NVLink bytes of the fp32 kernel above.
struct NvlsBf16Params {
  NvlsParams base;                 // mc / local address the bf16 buffer; n counts VALUES (multiple of 8)
  const float* in;                 // fp32 gradient (not inside the multicast buffer)
};

__device__ __forceinline__ void nvls_allreduce_bf16_body(const NvlsBf16Params& P) {
  const NvlsParams& p = P.base;
  const int nblocks = gridDim.x;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(p.epoch) + 1u;
  const int64_t n8 = p.n / 8;
  const int64_t per = (n8 + p.world - 1) / p.world;
  const int64_t lo = per * p.rank, hi = min(n8, lo + per);
  const int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  const int64_t nthreads = int64_t(nblocks) * blockDim.x;
  float lr = p.opt.lr, bc1 = 1.f, bc2 = 1.f;
  bool first_step = false;
  if (p.opt.kind) {
    if (p.opt.lr_ptr) lr = *p.opt.lr_ptr;
    const float t = p.opt.step_ptr ? *p.opt.step_ptr : 1.f;
    bc1 = 1.f - powf(p.opt.beta1, t);
    bc2 = 1.f - powf(p.opt.beta2, t);
    first_step = t <= 1.f;
  }
  uint4* local = reinterpret_cast<uint4*>(p.local);
  for (int64_t i = tid; i < n8; i += nthreads) {
    const float4 g0 = reinterpret_cast<const float4*>(P.in)[2 * i], g1 = reinterpret_cast<const float4*>(P.in)[2 * i + 1];
    local[i] = make_uint4(pack2_bf16(g0.x, g0.y), pack2_bf16(g0.z, g0.w), pack2_bf16(g1.x, g1.y), pack2_bf16(g1.z, g1.w));
  }
  flags_signal_and_wait(p.flags, p.rank, p.world, e, 0, p.counters + 0, nblocks);
  constexpr int kU = 4;
  uint4* mc = reinterpret_cast<uint4*>(p.mc);
  for (int64_t i0 = lo + tid; i0 < hi; i0 += nthreads * kU) {
    uint4 v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t i = i0 + u * nthreads;
      if (i < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(mc + i) : "memory");
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t i = i0 + u * nthreads;
      if (i < hi) {
        float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        acc_bf16x8(s, v[u]);
        const uint32_t r0 = pack2_bf16(s[0] * p.scale, s[1] * p.scale), r1 = pack2_bf16(s[2] * p.scale, s[3] * p.scale),
                       r2 = pack2_bf16(s[4] * p.scale, s[5] * p.scale), r3 = pack2_bf16(s[6] * p.scale, s[7] * p.scale);
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(mc + i), "f"(__uint_as_float(r0)), "f"(__uint_as_float(r1)), "f"(__uint_as_float(r2)),
                       "f"(__uint_as_float(r3)) : "memory");
      }
    }
  }
  flags_signal_and_wait(p.flags, p.rank, p.world, e, 1, p.counters + 2, nblocks);
  for (int64_t i = tid; i < n8; i += nthreads) {
    float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    acc_bf16x8(s, local[i]);
    apply_update(p.opt, i * 8, make_float4(s[0], s[1], s[2], s[3]), bc1, bc2, lr, first_step, p.out);
    apply_update(p.opt, i * 8 + 4, make_float4(s[4], s[5], s[6], s[7]), bc1, bc2, lr, first_step, p.out);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(p.counters + 1, 1u);
    if (prev == uint32_t(nblocks - 1)) {
      p.counters[1] = 0;
      *p.epoch = e;
    }
  }
}

__global__ void __launch_bounds__(kArThreads)
nvls_allreduce_bf16_kernel(const __grid_constant__ NvlsBf16Params p) { nvls_allreduce_bf16_body(p); }
__global__ void __launch_bounds__(kArOverlapThreads, 6)
nvls_allreduce_bf16_overlap_kernel(const __grid_constant__ NvlsBf16Params p) { nvls_allreduce_bf16_body(p); }

}  // namespace rdp

extern "C" int rdp_allreduce_nvls(void* mc_ptr, void* local_ptr, void* const* flag_ptrs, int rank, int world,
                                  uint32_t* epoch, uint32_t* counters, const rdp_opt_args* opt, float* out,
                                  int64_t n, float scale, int overlap_shape, void* stream) {
  using namespace rdp;
  RDP_REQUIRE(mc_ptr != nullptr, "rdp_allreduce_nvls: no multicast mapping (NVLS unsupported on this system)");
  RDP_REQUIRE(world >= 1 && world <= kMaxRanks && n > 0 && n % 4 == 0, "rdp_allreduce_nvls: bad world/n");
  NvlsParams p;
  memset(&p, 0, sizeof(p));
  p.mc = static_cast<float*>(mc_ptr);
  p.local = static_cast<float*>(local_ptr);
  for (int q = 0; q < world; ++q) p.flags[q] = static_cast<uint32_t*>(flag_ptrs[q]);
  p.rank = rank; p.world = world; p.n = n; p.scale = scale; p.epoch = epoch; p.counters = counters; p.out = out;
  if (opt != nullptr)
    p.opt = OptSpec{opt->kind, opt->p, opt->m, opt->v, opt->lr, opt->beta1, opt->beta2, opt->eps, opt->weight_decay,
                    opt->momentum, opt->dampening, opt->adamw, opt->nesterov, opt->lr_ptr, opt->step_ptr,
                    static_cast<__nv_bfloat16*>(opt->shadow)};
  int64_t work = n / 4 / world / 4;
  const int threads = overlap_shape ? kArOverlapThreads : kArThreads;
  int blocks = int((work + threads - 1) / threads);
  const int cap = num_sms();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (overlap_shape) nvls_allreduce_overlap_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  else nvls_allreduce_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// bf16-wire NVLS all-reduce: the multicast buffer (mc_ptr / local_ptr) holds n bf16 values; `in` is the fp32
// gradient; the averaged result goes through the optimizer epilogue and / or to `out` (fp32).  n % 8 == 0.
extern "C" int rdp_allreduce_nvls_bf16(void* mc_ptr, void* local_ptr, void* const* flag_ptrs, int rank, int world,
                                       uint32_t* epoch, uint32_t* counters, const rdp_opt_args* opt, const float* in,
                                       float* out, int64_t n, float scale, int overlap_shape, void* stream) {
  using namespace rdp;
  RDP_REQUIRE(mc_ptr != nullptr, "rdp_allreduce_nvls_bf16: no multicast mapping (NVLS unsupported on this system)");
  RDP_REQUIRE(world >= 1 && world <= kMaxRanks && n > 0 && n % 8 == 0, "rdp_allreduce_nvls_bf16: bad world/n");
  RDP_REQUIRE(in != nullptr && (opt != nullptr || out != nullptr), "rdp_allreduce_nvls_bf16: missing in / out");
  NvlsBf16Params P;
  memset(&P, 0, sizeof(P));
  NvlsParams& p = P.base;
  p.mc = static_cast<float*>(mc_ptr);
  p.local = static_cast<float*>(local_ptr);
  for (int q = 0; q < world; ++q) p.flags[q] = static_cast<uint32_t*>(flag_ptrs[q]);
  p.rank = rank; p.world = world; p.n = n; p.scale = scale; p.epoch = epoch; p.counters = counters; p.out = out;
  P.in = in;
  if (opt != nullptr)
    p.opt = OptSpec{opt->kind, opt->p, opt->m, opt->v, opt->lr, opt->beta1, opt->beta2, opt->eps, opt->weight_decay,
                    opt->momentum, opt->dampening, opt->adamw, opt->nesterov, opt->lr_ptr, opt->step_ptr,
                    static_cast<__nv_bfloat16*>(opt->shadow)};
  int64_t work = n / 8 / world / 4;
  const int threads = overlap_shape ? kArOverlapThreads : kArThreads;
  int blocks = int((work + threads - 1) / threads);
  const int cap = num_sms();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (overlap_shape) nvls_allreduce_bf16_overlap_kernel<<<blocks, threads, 0, s>>>(P);
  else nvls_allreduce_bf16_kernel<<<blocks, threads, 0, s>>>(P);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ================================================================================================
// All-to-all over peer memory (the standalone form of the K4 exchange)
// ================================================================================================
// Rank r stores chunk q of its input straight into slot r of rank q's receive buffer (NVSwitch gives
// every pair full bandwidth, so there is no schedule beyond a per-rank rotation of the start peer).
namespace rdp {

struct A2AParams {
  const float4* in;                // [world][chunk4]
  float* recv[kMaxRanks];          // recv[q]: rank q's receive buffer [world][chunk4 * 4 floats]
  uint32_t* flags[kMaxRanks];
  int rank, world;
  int64_t chunk4;                  // float4 vectors per (src, dst) pair
  int64_t parity_stride;           // floats between the two receive buffers (0: one buffer + a "consumed" round)
  uint32_t* epoch;
  uint32_t* counters;              // 16 words: [0] round A, [1] finished, [4 + q] blocks done with destination slot q
};

__global__ void __launch_bounds__(kArThreads)
alltoall_kernel(const __grid_constant__ A2AParams p) {
  const int nblocks = gridDim.x;
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(p.epoch) + 1u;
  // Receive buffers alternate on epoch parity (parity_stride != 0), which makes an "everybody has consumed the
  // last result" round unnecessary: this rank can only finish call e after every rank has ENTERED call e (it
  // needs their pushes), i.e. after their streams ran past the consumers of call e-1; so when it starts call
  // e+1 -- writing the buffers of call e-1 -- nobody reads them any more.  One flag round per call.
  const int64_t parity_off = int64_t(e & 1u) * p.parity_stride;
  if (p.parity_stride == 0)      // single buffer: round A (slot 3) says last call's receive buffers are consumed
    flags_signal_and_wait(p.flags, p.rank, p.world, e, 3, p.counters + 0, nblocks);
  // The grid is a multiple of `world` blocks: block b pushes to destination (rank + b % world) % world, so
  // all destinations are in flight at once (the switch gives every pair full bandwidth) and ranks start on
  // different peers.  The last block to finish a destination signals THAT destination; nobody waits for the
  // slowest destination before signalling the others.
  {
    const int q = blockIdx.x % p.world;
    const int sub = blockIdx.x / p.world, nsub = nblocks / p.world;
    const int dst = (p.rank + q) % p.world;
    const float4* src = p.in + int64_t(dst) * p.chunk4;
    float4* out = reinterpret_cast<float4*>(p.recv[dst] + parity_off) + int64_t(p.rank) * p.chunk4;
    const int64_t stride = int64_t(nsub) * blockDim.x;
    for (int64_t i0 = int64_t(sub) * blockDim.x + threadIdx.x; i0 < p.chunk4; i0 += 4 * stride) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (i0 + u * stride < p.chunk4) v[u] = src[i0 + u * stride];
#pragma unroll
      for (int u = 0; u < 4; ++u) if (i0 + u * stride < p.chunk4) out[i0 + u * stride] = v[u];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t prev = atomicAdd(p.counters + 4 + q, 1u);
      if (prev == uint32_t(nsub - 1)) {
        p.counters[4 + q] = 0;
        __threadfence_system();
        st_release_sys(p.flags[dst] + 4 * kMaxRanks + p.rank, e);
      }
    }
  }
  // round B (slot 4): the pushes of every rank have landed here
  if (int(threadIdx.x) < p.world) wait_epoch(p.flags[p.rank] + 4 * kMaxRanks + threadIdx.x, e, 0x304u);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(p.counters + 1, 1u);
    if (prev == uint32_t(nblocks - 1)) {
      p.counters[1] = 0;
      *p.epoch = e;
    }
  }
}

}  // namespace rdp

extern "C" int rdp_alltoall(const void* in, void* const* recv_bases, void* const* flag_ptrs, int rank, int world,
                            int64_t floats_per_pair, int64_t parity_stride_floats, uint32_t* epoch,
                            uint32_t* counters, void* stream) {
  using namespace rdp;
  RDP_REQUIRE(world >= 1 && world <= kMaxRanks && floats_per_pair > 0 && floats_per_pair % 4 == 0,
              "rdp_alltoall: bad world / chunk (%lld floats per pair must be a multiple of 4)", (long long)floats_per_pair);
  A2AParams p;
  memset(&p, 0, sizeof(p));
  p.in = static_cast<const float4*>(in);
  for (int q = 0; q < world; ++q) {
    p.recv[q] = static_cast<float*>(recv_bases[q]);
    p.flags[q] = static_cast<uint32_t*>(flag_ptrs[q]);
  }
  p.rank = rank; p.world = world; p.chunk4 = floats_per_pair / 4; p.epoch = epoch; p.counters = counters;
  p.parity_stride = parity_stride_floats;
  const int64_t work = p.chunk4 / 4;                       // 4 vectors per thread, per destination
  int per_dst = int((work + kArThreads - 1) / kArThreads);
  const int cap = num_sms() / world > 0 ? num_sms() / world : 1;
  if (per_dst > cap) per_dst = cap;
  if (per_dst < 1) per_dst = 1;
  const int blocks = per_dst * world;
  alltoall_kernel<<<blocks, kArThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  RDP_CHECK_CUDA(cudaGetLastError());
  return 0;
}
