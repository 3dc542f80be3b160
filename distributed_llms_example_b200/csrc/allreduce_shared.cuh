// Definitions shared by allreduce.cu and nvls.cu: flag-pad layout, cross-GPU barrier, wire
// formats, the communicator object.
#pragma once

#include <cuda.h>

#include <cstdlib>
#include <vector>

#include "common.cuh"

namespace b2 {

constexpr int kArThreads = 512;
constexpr int kArMaxBlocks = 128;                                        // rows of the flag pad
constexpr size_t kPadBytes = 16384;                                      // flags + status word
constexpr size_t kStatusOff = kArMaxBlocks * B2_MAX_WORLD * 4;           // after the flags
constexpr size_t kGateOff = kStatusOff + 256;                            // B2_MAX_WORLD gate flags
constexpr size_t kDoneOff = kGateOff + 256;                              // arrival counter of the sum-of-squares fold
constexpr size_t kAccOff = kDoneOff + 256;                               // double: sum of squares since the last take
constexpr size_t kSqOff = 8192;                                          // double[B2_MAX_WORLD][kArMaxBlocks] partials
static_assert(kAccOff + 8 <= kSqOff, "flag pad layout overlaps");
static_assert(kSqOff + B2_MAX_WORLD * kArMaxBlocks * 8 <= kPadBytes, "flag pad too small");

struct RankView {
  float* buf;                        // this rank's bucket, reduced in place
  char* stage[B2_MAX_WORLD];         // staging half of every rank, as mapped by this rank
  uint32_t* flags[B2_MAX_WORLD];     // flag pad of every rank, as mapped by this rank
  uint32_t* status;                  // this rank's sticky timeout word (device memory, in its pad)
  uint32_t* host_status;             // the same word mirrored into mapped pinned host memory
  int rank;
  int sumsq;                         // 0: off; k > 0: fold the sum of squares of the result, k ranks contribute
};

template <int NV>
struct ArArgs {
  RankView v[NV];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acquire_sys() {
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}
// Peer (or own) staging written by another GPU during this kernel: relaxed system-scope load,
// never the non-coherent path.
__device__ __forceinline__ uint4 ld_peer16(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// ---- a peer that does not arrive is an ERROR, never silently wrong gradients -------------------
// A rank that gives up at a barrier (B2_AR_TIMEOUT_MS) raises the sticky status word of EVERY
// rank (its own pad, its mapped host word, every peer's pad), and from then on every kernel of
// this communicator on every rank fills the bucket it was given with NaN instead of reducing
// it: the waiting rank never touches half-written peer data, the late rank cannot slip through
// on stale flags, and the host finds out with one plain load (b2_comm_poll), no synchronise.
template <int W>
__device__ __forceinline__ void raise_timeout(const RankView& v) {
  atomicExch(v.status, 1u);
#pragma unroll
  for (int k = 0; k < W; ++k)
    st_release_sys(reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(v.flags[k]) + kStatusOff), 1u);
  *reinterpret_cast<volatile uint32_t*>(v.host_status) = 1u;
  __threadfence_system();
}
__device__ __forceinline__ bool comm_poisoned(const RankView& v) { return ld_relaxed_sys(v.status) != 0u; }

// NaN-fill this CTA's share of the local bucket and mirror the status word to the host.
__device__ __forceinline__ void poison_local(const RankView& v, long long numel) {
  const float nan = __int_as_float(0x7fc00000);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
       i += (long long)gridDim.x * blockDim.x)
    v.buf[i] = nan;
  if (threadIdx.x == 0) {
    *reinterpret_cast<volatile uint32_t*>(v.host_status) = 1u;
    __threadfence_system();
  }
}

// All CTAs with the same blockIdx.x on all ranks meet here. Data written by any thread of
// the CTA before the call is visible to the peers' CTAs after it (bar.sync orders the CTA's
// writes before the releasing store; the peer's acquiring load orders its reads after).
// Returns false (to every thread of the CTA) if a peer did not arrive in time or the
// communicator is already poisoned: the caller must then poison its bucket and return.
template <int W>
__device__ __forceinline__ bool cross_barrier(const RankView& v, uint32_t epoch,
                                              unsigned long long timeout_ns) {
  __syncthreads();
  int bad = 0;
  if (W > 1 && threadIdx.x < W) {
    const int peer = threadIdx.x;
    st_release_sys(v.flags[peer] + blockIdx.x * B2_MAX_WORLD + v.rank, epoch);
    const uint32_t* mine = v.flags[v.rank] + blockIdx.x * B2_MAX_WORLD + peer;
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    // poll with relaxed loads (cheap), then one acquire fence once the flag has arrived
    while ((int)(ld_relaxed_sys(mine) - epoch) < 0) {
      if ((++spins & 1023u) == 0) {
        if (comm_poisoned(v)) { bad = 1; break; }
        if (globaltimer_ns() - t0 > timeout_ns) { bad = 2; break; }
      }
    }
    fence_acquire_sys();
  }
  bad = __syncthreads_or(bad);
  if (bad && threadIdx.x == 0) raise_timeout<W>(v);
  return bad == 0;
}

// ---- sum of squares of the reduced values, folded into the all-reduce ----
// The thread that produced a reduced pack adds its squares to a private fp32 accumulator;
// block_sumsq() folds the CTA in a fixed order and stores the partial into slot [rank][block] of
// EVERY rank's pad before the kernel's last barrier; fold_sumsq(), after that barrier, lets the
// last CTA of this rank to get there add the world x grid partials in a fixed order (double) onto
// the communicator's running total. Same partials, same order on every rank: bit-identical
// totals, deterministic run to run. Consumed by b2_comm_sumsq_clip.
__device__ __forceinline__ float sq4(uint4 o) {
  const float x = __uint_as_float(o.x), y = __uint_as_float(o.y), z = __uint_as_float(o.z), w = __uint_as_float(o.w);
  return __fmaf_rn(x, x, __fmaf_rn(y, y, __fmaf_rn(z, z, __fmul_rn(w, w))));
}
template <int W>
__device__ __forceinline__ void block_sumsq(const RankView& v, float acc, bool to_peers) {
  __shared__ double s_sq[kArThreads / 32];
  double d = (double)acc;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  if ((threadIdx.x & 31) == 0) s_sq[threadIdx.x >> 5] = d;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < kArThreads / 32; ++i) t += s_sq[i];
    const size_t slot = kSqOff + ((size_t)v.rank * kArMaxBlocks + blockIdx.x) * 8;
    if (to_peers) {
#pragma unroll
      for (int k = 0; k < W; ++k)
        *reinterpret_cast<volatile double*>(reinterpret_cast<char*>(v.flags[k]) + slot) = t;
    } else {
      *reinterpret_cast<volatile double*>(reinterpret_cast<char*>(v.flags[v.rank]) + slot) = t;
    }
    __threadfence_system();
  }
}
// Call after the barrier that follows block_sumsq, by all threads of every CTA. `first`/`count`
// select the rank slots that hold contributions (all W for the sliced algorithms, this rank's
// own slot for one-shot, where every rank reduces everything itself).
__device__ __forceinline__ void fold_sumsq(const RankView& v, int first, int count) {
  __shared__ bool s_last;
  char* pad = reinterpret_cast<char*>(v.flags[v.rank]);
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(reinterpret_cast<unsigned int*>(pad + kDoneOff), 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last || threadIdx.x >= 32) return;
  __threadfence();
  double t = 0.0;
  for (int k = first; k < first + count; ++k)
    for (int b = threadIdx.x; b < (int)gridDim.x; b += 32)
      t += *reinterpret_cast<const volatile double*>(pad + kSqOff + ((size_t)k * kArMaxBlocks + b) * 8);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if (threadIdx.x == 0) {
    double* acc = reinterpret_cast<double*>(pad + kAccOff);
    *acc += t;
    *reinterpret_cast<unsigned int*>(pad + kDoneOff) = 0u;
  }
}

// One tiny CTA per rank that meets its peers BEFORE the wide kernel is launched behind it on the
// same stream: ranks reach a bucket at different times (the training step is host-bound more
// often than not), and waiting for a late peer should cost one SM, not the 64 the data
// kernel occupies. After the gate every rank's bucket is complete and the data kernels start
// within launch skew of each other.
template <bool BF16W>
struct Wire;
template <>
struct Wire<false> {
  static constexpr int kElems = 4;  // elements per 16-byte staged pack
  static __device__ __forceinline__ uint4 load_scaled(const float* buf, long long pack, float scale) {
    uint4 u = ld_stream16(buf + pack * 4);
    u.x = __float_as_uint(__fmul_rn(__uint_as_float(u.x), scale));
    u.y = __float_as_uint(__fmul_rn(__uint_as_float(u.y), scale));
    u.z = __float_as_uint(__fmul_rn(__uint_as_float(u.z), scale));
    u.w = __float_as_uint(__fmul_rn(__uint_as_float(u.w), scale));
    return u;
  }
  static __device__ __forceinline__ void to_f(uint4 u, float* x) {
    x[0] = __uint_as_float(u.x); x[1] = __uint_as_float(u.y);
    x[2] = __uint_as_float(u.z); x[3] = __uint_as_float(u.w);
  }
  static __device__ __forceinline__ uint4 from_f(const float* x) {
    return make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]), __float_as_uint(x[2]),
                      __float_as_uint(x[3]));
  }
  static __device__ __forceinline__ void store_buf(float* buf, long long pack, uint4 staged) {
    *reinterpret_cast<uint4*>(buf + pack * 4) = staged;
  }
  static __device__ __forceinline__ float stage1(const float* buf, long long e, float scale) {
    return __fmul_rn(buf[e], scale);
  }
  // tail elements (numel % kElems) live after the last whole pack of the staging half
  static __device__ __forceinline__ void tail_store(char* base, int i, float f) {
    reinterpret_cast<float*>(base)[i] = f;
  }
  static __device__ __forceinline__ float tail_load(const char* base, int i) {
    return reinterpret_cast<const volatile float*>(base)[i];
  }
  static __device__ __forceinline__ float round_wire(float f) { return f; }
  static __device__ __forceinline__ float sq(uint4 staged) { return sq4(staged); }  // sum of squares of the pack
};
template <>
struct Wire<true> {
  static constexpr int kElems = 8;
  static __device__ __forceinline__ uint4 load_scaled(const float* buf, long long pack, float scale) {
    const uint4 a = ld_stream16(buf + pack * 8), b = ld_stream16(buf + pack * 8 + 4);
    return make_uint4(
        pack_bf16(__uint_as_float(a.x) * scale, __uint_as_float(a.y) * scale),
        pack_bf16(__uint_as_float(a.z) * scale, __uint_as_float(a.w) * scale),
        pack_bf16(__uint_as_float(b.x) * scale, __uint_as_float(b.y) * scale),
        pack_bf16(__uint_as_float(b.z) * scale, __uint_as_float(b.w) * scale));
  }
  static __device__ __forceinline__ void to_f(uint4 u, float* x) {
    x[0] = bf16lo(u.x); x[1] = bf16hi(u.x); x[2] = bf16lo(u.y); x[3] = bf16hi(u.y);
    x[4] = bf16lo(u.z); x[5] = bf16hi(u.z); x[6] = bf16lo(u.w); x[7] = bf16hi(u.w);
  }
  static __device__ __forceinline__ uint4 from_f(const float* x) {
    return make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]),
                      pack_bf16(x[6], x[7]));
  }
  static __device__ __forceinline__ void store_buf(float* buf, long long pack, uint4 staged) {
    float x[8];
    to_f(staged, x);
    uint4* o = reinterpret_cast<uint4*>(buf + pack * 8);
    o[0] = make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]), __float_as_uint(x[2]), __float_as_uint(x[3]));
    o[1] = make_uint4(__float_as_uint(x[4]), __float_as_uint(x[5]), __float_as_uint(x[6]), __float_as_uint(x[7]));
  }
  static __device__ __forceinline__ float stage1(const float* buf, long long e, float scale) {
    return round_wire(buf[e] * scale);
  }
  static __device__ __forceinline__ void tail_store(char* base, int i, float f) {
    reinterpret_cast<unsigned short*>(base)[i] = (unsigned short)(__float_as_uint(round_wire(f)) >> 16);
  }
  static __device__ __forceinline__ float tail_load(const char* base, int i) {
    return __uint_as_float((uint32_t) reinterpret_cast<const volatile unsigned short*>(base)[i] << 16);
  }
  static __device__ __forceinline__ float round_wire(float f) {
    return __bfloat162float(__float2bfloat16_rn(f));
  }
  static __device__ __forceinline__ float sq(uint4 staged) {
    float x[8], a = 0.f;
    to_f(staged, x);
#pragma unroll
    for (int e = 0; e < 8; ++e) a = __fmaf_rn(x[e], x[e], a);
    return a;
  }
};

// Sum of pack `i` over all ranks, rank order 0..W-1, fp32 adds (no fma, no reassociation).
static inline long long env_ll(const char* name, long long dflt) {
  const char* s = std::getenv(name);
  if (!s || !*s) return dflt;
  char* end = nullptr;
  long long v = std::strtoll(s, &end, 10);
  return (end && *end == 0) ? v : dflt;
}

// One-shot moves (W-1)*S inbound per rank and needs one barrier; two-shot moves
// 2(W-1)/W*S and needs two. At W=2 the bytes are equal, so one-shot always wins.
static inline int pick_blocks(long long npack, int max_blocks, long long packs_per_thread = 4) {
  // `packs_per_thread` 16-byte packs per thread before adding a CTA (B2_AR_PACKS_PER_THREAD,
  // default 4: a 25 MiB bucket at world 8 fills ~100 CTAs); small buckets stay on few CTAs so the
  // barrier is cheap and the backward GEMMs keep their SMs.
  const long long ppt = packs_per_thread < 1 ? 1 : packs_per_thread;
  long long nb = (npack + (long long)kArThreads * ppt - 1) / ((long long)kArThreads * ppt);
  if (nb < 1) nb = 1;
  if (nb > max_blocks) nb = max_blocks;
  return (int)nb;
}

}  // namespace b2

struct b2_nvls {  // NVLink multicast (in-switch reduction) state, see nvls.cu
  bool have_mc = false, ready = false;
  size_t bytes = 0;
  size_t stage_bytes = 0;  // [0, stage_bytes) is the staging area of b2_allreduce_nvls; the pool starts after it
  CUmemGenericAllocationHandle mc = 0, mem = 0;
  CUdeviceptr uc_va = 0, mc_va = 0;
};

struct b2_opened_handle {  // one cudaIpcOpenMemHandle per distinct peer allocation
  int peer;
  unsigned char handle[B2_IPC_HANDLE_BYTES];
  char* base;
};

struct b2_registration {
  bool live = false;
  size_t bytes = 0;
  char* ptr[B2_MAX_WORLD] = {nullptr};  // the registered buffer of every rank, as mapped here
};

struct b2_comm {
  b2_nvls nvls;
  std::vector<b2_opened_handle> peer_allocs;
  std::vector<b2_registration> regs;
  int rank = 0, world = 1, device = 0;
  size_t half_bytes = 0;
  char* base[B2_MAX_WORLD] = {nullptr};  // mapped symmetric allocation of every rank
  bool opened = false;
  uint32_t epoch = 1;
  int parity = 0;
  int max_blocks = 128;
  int unroll = 0;                       // B2_AR_UNROLL: loads-in-flight variant of the zero-copy kernels
  int ppt = 4;                          // B2_AR_PACKS_PER_THREAD: 16-byte packs per thread before adding a CTA
  unsigned long long timeout_ns = 600ull * 1000000000ull;
  uint32_t* host_status = nullptr;      // mapped pinned host word the kernels mirror the status into
  uint32_t* host_status_dev = nullptr;  // its device-side address
  bool sumsq = false;                   // fold the sum of squares of every reduced bucket
  long long sumsq_numel = 0;            // elements folded since the last b2_comm_sumsq_take
};

// Fills the parts of a RankView that do not depend on the buffers of the call.
static inline void fill_view_common(const b2_comm* c, b2::RankView& v) {
  v.rank = c->rank;
  for (int r = 0; r < B2_MAX_WORLD; ++r) v.flags[r] = (uint32_t*)c->base[r < c->world ? r : c->rank];
  v.status = (uint32_t*)(c->base[c->rank] + b2::kStatusOff);
  v.host_status = c->host_status_dev;
  v.sumsq = 0;
}


int b2_nvls_release(b2_comm* c);            // nvls.cu
const void* b2_gate_kernel_for(int world);  // allreduce.cu: the one-CTA gate kernel for a world size
