// Gradient mean all-reduce over NVLink peer memory (include/b2ddp.h, section (a)).
//
// Replaces ncclAllReduce as driven by the DDP Reducer (SP/torch/nn/parallel/distributed.py
// :1255-1284 -> default_comm_hooks.hpp:36-51; reference call sites train-accelerator.py:189,222
// and train-torchrun.py:127,140) and by train-task.py:65-69 (per-tensor all_reduce + "/= size").
//
// One process per GPU. Every rank owns a "symmetric" allocation [flag pad | staging half 0 |
// staging half 1] that all peers map through CUDA IPC; NVSwitch gives each GPU full bandwidth
// to every peer, so there is no ring: every CTA talks to all peers at once.
//
//   one kernel per call:
//     copy-in   bucket (ordinary caching-allocator memory) -> own staging, fused with the
//               1/world scale and the optional fp32->bf16 wire cast
//     barrier   per-CTA flags in peer memory (st.release.sys / ld.acquire.sys)
//     one-shot  every rank pulls all W staged copies, adds them in rank order 0..W-1 in fp32
//               and writes the bucket                     (latency-optimal, (W-1)*S inbound)
//     two-shot  rank r reduces slice r the same way, publishes it in its own staging,
//               barrier, every rank pulls the other W-1 reduced slices
//                                                  (bandwidth-optimal, 2(W-1)/W*S inbound)
//
// Fixed rank-order summation makes the result bit-identical on every rank for both
// algorithms. Consecutive calls alternate staging halves, which is what makes a trailing
// barrier unnecessary: a rank can only be two calls ahead of a peer's reads if it passed
// the barrier of the call in between, which that peer enters only after finishing its reads.
// Calls on one b2_comm must therefore be stream-ordered on every rank.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "allreduce_shared.cuh"

namespace b2 {

template <int W>
__global__ void __launch_bounds__(32, 1)
gate_kernel(const __grid_constant__ ArArgs<1> args, uint32_t epoch, unsigned long long timeout_ns) {
  const RankView& v = args.v[0];
  int bad = 0;
  if (threadIdx.x < W) {
    const int peer = threadIdx.x;
    uint32_t* gate_peer = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(v.flags[peer]) + kGateOff);
    const uint32_t* gate_mine = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(v.flags[v.rank]) + kGateOff);
    st_release_sys(gate_peer + v.rank, epoch);
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    while ((int)(ld_relaxed_sys(gate_mine + peer) - epoch) < 0) {
      __nanosleep(100);
      if ((++spins & 255u) == 0) {
        if (comm_poisoned(v)) { bad = 1; break; }
        if (globaltimer_ns() - t0 > timeout_ns) { bad = 2; break; }
      }
    }
    fence_acquire_sys();
  }
  // the data kernel behind this gate sees the raised status word and poisons its bucket
  if (__any_sync(0xffffffffu, bad) && threadIdx.x == 0) raise_timeout<W>(v);
}

template <int W, bool BF16W>
__device__ __forceinline__ void reduce_packs(const RankView& v, long long i, float* acc) {
  using Wr = Wire<BF16W>;
  uint4 r[W];
#pragma unroll
  for (int k = 0; k < W; ++k) r[k] = ld_peer16(v.stage[k] + i * 16);
  Wr::to_f(r[0], acc);
#pragma unroll
  for (int k = 1; k < W; ++k) {
    float x[Wr::kElems];
    Wr::to_f(r[k], x);
#pragma unroll
    for (int e = 0; e < Wr::kElems; ++e) acc[e] = __fadd_rn(acc[e], x[e]);
  }
}

template <int W, bool BF16W, int ALGO, int NV>
__global__ void __launch_bounds__(kArThreads, 1)
allreduce_kernel(const __grid_constant__ ArArgs<NV> args, long long numel, float scale,
                 uint32_t epoch, unsigned long long timeout_ns) {
  using Wr = Wire<BF16W>;
  constexpr int EPP = Wr::kElems;
  constexpr bool kOneShot = (ALGO == B2_ALGO_ONE_SHOT || W == 1);
  const RankView& v = args.v[NV == 1 ? 0 : blockIdx.y];
  const int rank = v.rank;
  const int nb = gridDim.x, b = blockIdx.x, tid = threadIdx.x;
  const long long npack = numel / EPP;
  float* __restrict__ buf = v.buf;
  char* own = v.stage[rank];
  // Elements past the last whole pack ride in block 0, one-shot style, for both algorithms.
  const long long tail0 = npack * EPP;
  char* own_tail = own + npack * 16;
  float sq = 0.f;  // squares of the reduced values this thread produced (v.sumsq)

  if (comm_poisoned(v)) { poison_local(v, numel); return; }
  if (kOneShot) {
    const long long c = (npack + nb - 1) / nb;
    const long long lo = min((long long)b * c, npack), hi = min(lo + c, npack);
    {
      constexpr int UC = 4;  // copy-in: 4 independent 16-byte (fp32) / 32-byte (bf16 wire) reads in flight
      long long i = lo + tid;
      for (; i + (UC - 1) * kArThreads < hi; i += UC * kArThreads) {
        uint4 t[UC];
#pragma unroll
        for (int u = 0; u < UC; ++u) t[u] = Wr::load_scaled(buf, i + u * kArThreads, scale);
#pragma unroll
        for (int u = 0; u < UC; ++u) *reinterpret_cast<uint4*>(own + (i + u * kArThreads) * 16) = t[u];
      }
      for (; i < hi; i += kArThreads)
        *reinterpret_cast<uint4*>(own + i * 16) = Wr::load_scaled(buf, i, scale);
    }
    if (b == 0 && tid < numel - tail0) Wr::tail_store(own_tail, tid, Wr::stage1(buf, tail0 + tid, scale));
    if (!cross_barrier<W>(v, epoch, timeout_ns)) { poison_local(v, numel); return; }
    constexpr int U = (W >= 8) ? 2 : 4;
    long long i = lo + tid;
    for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
      float acc[U][EPP];
#pragma unroll
      for (int u = 0; u < U; ++u) reduce_packs<W, BF16W>(v, i + u * kArThreads, acc[u]);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint4 o = Wr::from_f(acc[u]);
        sq += Wr::sq(o);
        Wr::store_buf(buf, i + u * kArThreads, o);
      }
    }
    for (; i < hi; i += kArThreads) {
      float acc[EPP];
      reduce_packs<W, BF16W>(v, i, acc);
      const uint4 o = Wr::from_f(acc);
      sq += Wr::sq(o);
      Wr::store_buf(buf, i, o);
    }
  } else {
    const long long Q = (npack + W - 1) / W;   // packs per rank slice
    const long long q = (Q + nb - 1) / nb;     // packs per (slice, block) sub-range
    // copy-in: this CTA stages its sub-range of EVERY slice, because the peer CTA with the
    // same blockIdx is the one that will read it after the per-CTA barrier.
#pragma unroll 1
    for (int s = 0; s < W; ++s) {
      const long long slo = min((long long)s * Q, npack), shi = min(slo + Q, npack);
      const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
      constexpr int UC = 4;
      long long i = lo + tid;
      for (; i + (UC - 1) * kArThreads < hi; i += UC * kArThreads) {
        uint4 t[UC];
#pragma unroll
        for (int u = 0; u < UC; ++u) t[u] = Wr::load_scaled(buf, i + u * kArThreads, scale);
#pragma unroll
        for (int u = 0; u < UC; ++u) *reinterpret_cast<uint4*>(own + (i + u * kArThreads) * 16) = t[u];
      }
      for (; i < hi; i += kArThreads)
        *reinterpret_cast<uint4*>(own + i * 16) = Wr::load_scaled(buf, i, scale);
    }
    if (b == 0 && tid < numel - tail0) Wr::tail_store(own_tail, tid, Wr::stage1(buf, tail0 + tid, scale));
    if (!cross_barrier<W>(v, epoch, timeout_ns)) { poison_local(v, numel); return; }
    // reduce-scatter: my slice, my sub-range, summed over all ranks; publish in place.
    {
      const long long slo = min((long long)rank * Q, npack), shi = min(slo + Q, npack);
      const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
      constexpr int U = (W >= 8) ? 2 : 4;
      long long i = lo + tid;
      for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
        float acc[U][EPP];
#pragma unroll
        for (int u = 0; u < U; ++u) reduce_packs<W, BF16W>(v, i + u * kArThreads, acc[u]);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint4 o = Wr::from_f(acc[u]);
          sq += Wr::sq(o);
          *reinterpret_cast<uint4*>(own + (i + u * kArThreads) * 16) = o;
          Wr::store_buf(buf, i + u * kArThreads, o);
        }
      }
      for (; i < hi; i += kArThreads) {
        float acc[EPP];
        reduce_packs<W, BF16W>(v, i, acc);
        const uint4 o = Wr::from_f(acc);
        sq += Wr::sq(o);
        *reinterpret_cast<uint4*>(own + i * 16) = o;
        Wr::store_buf(buf, i, o);
      }
    }
  }
  // tail elements (fewer than one pack): every rank sums them itself after the first barrier;
  // their squares are counted once (one-shot: by every rank for itself; two-shot: by rank 0)
  if (b == 0 && tid < numel - tail0) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < W; ++k) {
      const float x = Wr::tail_load(v.stage[k] + npack * 16, tid);
      acc = (k == 0) ? x : __fadd_rn(acc, x);
    }
    acc = Wr::round_wire(acc);
    buf[tail0 + tid] = acc;
    if (kOneShot || rank == 0) sq = __fmaf_rn(acc, acc, sq);
  }
  if (kOneShot) {
    if (v.sumsq) {
      block_sumsq<W>(v, sq, false);
      fold_sumsq(v, rank, 1);
    }
    return;
  }
  if (v.sumsq) block_sumsq<W>(v, sq, true);
  if (!cross_barrier<W>(v, epoch + 1, timeout_ns)) { poison_local(v, numel); return; }
  {
    const long long Q = (npack + W - 1) / W;
    const long long q = (Q + nb - 1) / nb;
    // all-gather: pull the other ranks' reduced sub-ranges, starting with my right-hand
    // neighbour so that the W ranks do not all hit the same peer at the same time.
    constexpr int U2 = 4;
#pragma unroll 1
    for (int k = 1; k < W; ++k) {
      const int s = (rank + k) % W;
      const char* src = v.stage[s];
      const long long slo = min((long long)s * Q, npack), shi = min(slo + Q, npack);
      const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
      long long i = lo + tid;
      for (; i + (U2 - 1) * kArThreads < hi; i += U2 * kArThreads) {
        uint4 r[U2];
#pragma unroll
        for (int u = 0; u < U2; ++u) r[u] = ld_peer16(src + (i + u * kArThreads) * 16);
#pragma unroll
        for (int u = 0; u < U2; ++u) Wr::store_buf(buf, i + u * kArThreads, r[u]);
      }
      for (; i < hi; i += kArThreads) Wr::store_buf(buf, i, ld_peer16(src + i * 16));
    }
  }
  if (v.sumsq) fold_sumsq(v, 0, W);
}


// ---- zero-copy variant: buckets registered for peer access -------------------------------
//
// When the bucket itself is mapped by every peer (b2_comm_register), the staging copy
// disappears: two-shot runs IN PLACE on the buckets.
//   barrier A   every rank's bucket holds its local gradients (its producers are done)
//   stage 1     rank r: slice r <- sum_k scale * bucket_k[slice r], rank order, written into
//               its OWN bucket (only rank r ever reads slice r of rank r's bucket in stage 1)
//   barrier B   every slice owner has published its reduced slice
//   stage 2     rank r pulls reduced slice k from bucket_k for k != r into its own bucket
//               (the CTA of rank k that read this sub-range in stage 1 is past barrier B)
//   barrier C   nobody is still reading this rank's reduced slice when the kernel returns and
//               the next backward starts overwriting the bucket
// stage[k] of the RankView holds bucket_k as mapped by this rank; buf is bucket_rank.
template <int W, int NV>
__global__ void __launch_bounds__(kArThreads, 1)
allreduce_inplace_kernel(const __grid_constant__ ArArgs<NV> args, long long numel, float scale,
                         uint32_t epoch, unsigned long long timeout_ns, int gated) {
  const RankView& v = args.v[NV == 1 ? 0 : blockIdx.y];
  const int rank = v.rank;
  const int nb = gridDim.x, b = blockIdx.x, tid = threadIdx.x;
  const long long npack = numel / 4;
  const long long tail0 = npack * 4;
  float* __restrict__ buf = v.buf;
  const long long Q = (npack + W - 1) / W;
  const long long q = (Q + nb - 1) / nb;
  constexpr int U = (W >= 8) ? 2 : 4;
  float sq = 0.f;

  if (comm_poisoned(v)) { poison_local(v, numel); return; }
  if (!gated && !cross_barrier<W>(v, epoch, timeout_ns)) {  // barrier A, unless a gate_kernel ran first
    poison_local(v, numel);
    return;
  }
  {
    const long long slo = min((long long)rank * Q, npack), shi = min(slo + Q, npack);
    const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
    long long i = lo + tid;
    for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
      uint4 r[U][W];
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int k = 0; k < W; ++k) r[u][k] = ld_peer16(v.stage[k] + (i + u * kArThreads) * 16);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float acc[4];
#pragma unroll
        for (int k = 0; k < W; ++k) {
          float x[4];
          Wire<false>::to_f(r[u][k], x);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float t = __fmul_rn(x[e], scale);
            acc[e] = (k == 0) ? t : __fadd_rn(acc[e], t);
          }
        }
        const uint4 o = Wire<false>::from_f(acc);
        sq += sq4(o);
        *reinterpret_cast<uint4*>(buf + (i + u * kArThreads) * 4) = o;
      }
    }
    for (; i < hi; i += kArThreads) {
      float acc[4];
#pragma unroll
      for (int k = 0; k < W; ++k) {
        float x[4];
        Wire<false>::to_f(ld_peer16(v.stage[k] + i * 16), x);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float t = __fmul_rn(x[e], scale);
          acc[e] = (k == 0) ? t : __fadd_rn(acc[e], t);
        }
      }
      const uint4 o = Wire<false>::from_f(acc);
      sq += sq4(o);
      *reinterpret_cast<uint4*>(buf + i * 4) = o;
    }
  }
  // tail (numel % 4): every rank sums it redundantly now, writes it after barrier B
  float tail_acc = 0.f;
  const bool has_tail = (b == 0 && tid < numel - tail0);
  if (has_tail) {
#pragma unroll
    for (int k = 0; k < W; ++k) {
      const float t = __fmul_rn(Wire<false>::tail_load(v.stage[k] + tail0 * 4, tid), scale);
      tail_acc = (k == 0) ? t : __fadd_rn(tail_acc, t);
    }
    if (rank == 0) sq = __fmaf_rn(tail_acc, tail_acc, sq);
  }
  if (v.sumsq) block_sumsq<W>(v, sq, true);
  if (!cross_barrier<W>(v, epoch + 1, timeout_ns)) { poison_local(v, numel); return; }
  if (has_tail) buf[tail0 + tid] = tail_acc;
  constexpr int U2 = 4;
#pragma unroll 1
  for (int k = 1; k < W; ++k) {
    const int s = (rank + k) % W;
    const char* src = v.stage[s];
    const long long slo = min((long long)s * Q, npack), shi = min(slo + Q, npack);
    const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
    long long i = lo + tid;
    for (; i + (U2 - 1) * kArThreads < hi; i += U2 * kArThreads) {
      uint4 r[U2];
#pragma unroll
      for (int u = 0; u < U2; ++u) r[u] = ld_peer16(src + (i + u * kArThreads) * 16);
#pragma unroll
      for (int u = 0; u < U2; ++u) *reinterpret_cast<uint4*>(buf + (i + u * kArThreads) * 4) = r[u];
    }
    for (; i < hi; i += kArThreads) *reinterpret_cast<uint4*>(buf + i * 4) = ld_peer16(src + i * 16);
  }
  if (!cross_barrier<W>(v, epoch + 2, timeout_ns)) { poison_local(v, numel); return; }
  if (v.sumsq) fold_sumsq(v, 0, W);
}

// ---- zero-copy, one pass: reduce my slice from every bucket, PUSH the result into every bucket --
//
// With the peers' buckets mapped, the slice owner can also WRITE the reduced values straight
// into each peer's bucket. Reads arrive on the inbound half of the NVLink port while the
// stores leave on the outbound half, so the two halves of two-shot overlap instead of following
// each other, and the middle barrier disappears:
//   gate / barrier A   every rank's bucket is complete
//   one pass           rank r: for i in slice r: x = sum_k scale*bucket_k[i] (rank order);
//                      bucket_k[i] = x for every k            ((W-1)/W*S in, (W-1)/W*S out)
//   barrier C          all pushed values have landed everywhere
// Only rank r touches slice r of any bucket during the pass (it reads an element from every
// bucket before it overwrites that element), so there is no hazard inside the pass.
// Weak (fire-and-forget) 16-byte store: completion and visibility are established once, by the
// __threadfence_system() + releasing flag store of the final barrier, not per store.
// The owner holds every reduced element in registers, so the sum of squares the gradient clip
// needs is folded into this pass: no separate read of the gradients.
__device__ __forceinline__ void st_peer16(void* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

template <int W, int U>
__device__ __forceinline__ void push_range(const RankView& v, long long lo, long long hi, float scale, float& sq) {
  const int rank = v.rank, tid = threadIdx.x;
  long long i = lo + tid;
  for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
    uint4 r[U][W];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int k = 0; k < W; ++k) r[u][k] = ld_stream16(v.stage[k] + (i + u * kArThreads) * 16);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float acc[4];
#pragma unroll
      for (int k = 0; k < W; ++k) {
        float x[4];
        Wire<false>::to_f(r[u][k], x);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float t = __fmul_rn(x[e], scale);
          acc[e] = (k == 0) ? t : __fadd_rn(acc[e], t);
        }
      }
      const uint4 o = Wire<false>::from_f(acc);
      sq += sq4(o);
#pragma unroll
      for (int k = 0; k < W; ++k) st_peer16(v.stage[(rank + k) % W] + (i + u * kArThreads) * 16, o);
    }
  }
  for (; i < hi; i += kArThreads) {
    float acc[4];
#pragma unroll
    for (int k = 0; k < W; ++k) {
      float x[4];
      Wire<false>::to_f(ld_stream16(v.stage[k] + i * 16), x);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float t = __fmul_rn(x[e], scale);
        acc[e] = (k == 0) ? t : __fadd_rn(acc[e], t);
      }
    }
    const uint4 o = Wire<false>::from_f(acc);
    sq += sq4(o);
#pragma unroll
    for (int k = 0; k < W; ++k) st_peer16(v.stage[(rank + k) % W] + i * 16, o);
  }
}

template <int W, int NV>
__global__ void __launch_bounds__(kArThreads, 1)
allreduce_push_kernel(const __grid_constant__ ArArgs<NV> args, long long numel, float scale,
                      uint32_t epoch, unsigned long long timeout_ns, int gated, int unroll) {
  const RankView& v = args.v[NV == 1 ? 0 : blockIdx.y];
  const int rank = v.rank;
  const int nb = gridDim.x, b = blockIdx.x, tid = threadIdx.x;
  const long long npack = numel / 4;
  const long long tail0 = npack * 4;
  const long long Q = (npack + W - 1) / W;
  const long long q = (Q + nb - 1) / nb;
  float sq = 0.f;

  if (comm_poisoned(v)) { poison_local(v, numel); return; }
  if (!gated && !cross_barrier<W>(v, epoch, timeout_ns)) { poison_local(v, numel); return; }
  const long long slo = min((long long)rank * Q, npack), shi = min(slo + Q, npack);
  const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
  // loads in flight per thread = U x W x 16 B; B2_AR_UNROLL picks 0: the tuned default,
  // 1: 16/W packs, 2: half the default
  constexpr int kU0 = (W >= 8) ? 2 : 4, kU1 = (W >= 8) ? 2 : 16 / (W > 1 ? W : 2), kU2 = kU0 / 2;
  if (unroll == 1) push_range<W, kU1>(v, lo, hi, scale, sq);
  else if (unroll == 2) push_range<W, kU2>(v, lo, hi, scale, sq);
  else push_range<W, kU0>(v, lo, hi, scale, sq);
  // tail (numel % 4): rank 0 owns it
  if (rank == 0 && b == 0 && tid < numel - tail0) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < W; ++k) {
      const float t = __fmul_rn(Wire<false>::tail_load(v.stage[k] + tail0 * 4, tid), scale);
      acc = (k == 0) ? t : __fadd_rn(acc, t);
    }
    sq = __fmaf_rn(acc, acc, sq);
#pragma unroll
    for (int k = 0; k < W; ++k) reinterpret_cast<volatile float*>(v.stage[k] + tail0 * 4)[tid] = acc;
  }
  if (v.sumsq) block_sumsq<W>(v, sq, true);
  __threadfence_system();  // this thread's remote stores are performed before the CTA signals
  if (!cross_barrier<W>(v, epoch + 1, timeout_ns)) { poison_local(v, numel); return; }
  if (v.sumsq) fold_sumsq(v, 0, W);
}

// ---- host side ----------------------------------------------------------------------

using KernelFn = const void*;

template <int NV>
static KernelFn pick_kernel(int world, bool bf16w, int algo) {
#define B2_AR_CASE(Wv)                                                                        \
  if (world == Wv) {                                                                          \
    if (algo == B2_ALGO_ONE_SHOT)                                                             \
      return bf16w ? (KernelFn)allreduce_kernel<Wv, true, B2_ALGO_ONE_SHOT, NV>               \
                   : (KernelFn)allreduce_kernel<Wv, false, B2_ALGO_ONE_SHOT, NV>;             \
    return bf16w ? (KernelFn)allreduce_kernel<Wv, true, B2_ALGO_TWO_SHOT, NV>                 \
                 : (KernelFn)allreduce_kernel<Wv, false, B2_ALGO_TWO_SHOT, NV>;               \
  }
  B2_AR_CASE(1)
  B2_AR_CASE(2)
  B2_AR_CASE(4)
  B2_AR_CASE(8)
#undef B2_AR_CASE
  return nullptr;
}

template <int NV>
static KernelFn pick_push_kernel(int world) {
  switch (world) {
    case 1: return (KernelFn)allreduce_push_kernel<1, NV>;
    case 2: return (KernelFn)allreduce_push_kernel<2, NV>;
    case 4: return (KernelFn)allreduce_push_kernel<4, NV>;
    case 8: return (KernelFn)allreduce_push_kernel<8, NV>;
  }
  return nullptr;
}

template <int NV>
static KernelFn pick_inplace_kernel(int world) {
  switch (world) {
    case 1: return (KernelFn)allreduce_inplace_kernel<1, NV>;
    case 2: return (KernelFn)allreduce_inplace_kernel<2, NV>;
    case 4: return (KernelFn)allreduce_inplace_kernel<4, NV>;
    case 8: return (KernelFn)allreduce_inplace_kernel<8, NV>;
  }
  return nullptr;
}

static int pick_algo(int world, long long bytes) {
  if (world <= 2) return B2_ALGO_ONE_SHOT;
  const long long thresh = env_ll("B2_AR_ONE_SHOT_MAX_BYTES", world == 4 ? (4ll << 20) : (1ll << 20));
  return bytes <= thresh ? B2_ALGO_ONE_SHOT : B2_ALGO_TWO_SHOT;
}

}  // namespace b2

const void* b2_gate_kernel_for(int world) {
  return world == 2 ? (const void*)b2::gate_kernel<2>
       : world == 4 ? (const void*)b2::gate_kernel<4> : (const void*)b2::gate_kernel<8>;
}

extern "C" int b2_comm_create(int rank, int world, size_t staging_bytes, b2_comm** out) {
  using namespace b2;
  B2_REQUIRE(out, "comm_create: null out");
  B2_REQUIRE(world == 1 || world == 2 || world == 4 || world == 8,
             "comm_create: world %d not in {1,2,4,8}", world);
  B2_REQUIRE(rank >= 0 && rank < world, "comm_create: rank %d outside [0,%d)", rank, world);
  B2_REQUIRE(staging_bytes >= 4096, "comm_create: staging_bytes %zu < 4096", staging_bytes);
  b2_comm* c = new b2_comm();
  c->rank = rank;
  c->world = world;
  c->half_bytes = (staging_bytes + 255) & ~(size_t)255;
  c->max_blocks = (int)env_ll("B2_AR_BLOCKS", 128);
  if (c->max_blocks < 1) c->max_blocks = 1;
  if (c->max_blocks > kArMaxBlocks) c->max_blocks = kArMaxBlocks;
  c->unroll = (int)env_ll("B2_AR_UNROLL", 0);
  c->ppt = (int)env_ll("B2_AR_PACKS_PER_THREAD", 4);
  // A peer that has not arrived after this long is treated as dead (NCCL-watchdog scale: a
  // checkpoint save or an evaluation pass on one rank must not trip it).
  c->timeout_ns = (unsigned long long)env_ll("B2_AR_TIMEOUT_MS", 600000) * 1000000ull;
  cudaError_t e = cudaGetDevice(&c->device);
  if (e == cudaSuccess) e = cudaMalloc((void**)&c->base[rank], kPadBytes + 2 * c->half_bytes);
  if (e == cudaSuccess) e = cudaMemset(c->base[rank], 0, kPadBytes);
  // the status word's host mirror: mapped pinned memory, so b2_comm_poll is one host load
  if (e == cudaSuccess) e = cudaHostAlloc((void**)&c->host_status, 64, cudaHostAllocMapped);
  if (e == cudaSuccess) {
    *c->host_status = 0;
    e = cudaHostGetDevicePointer((void**)&c->host_status_dev, c->host_status, 0);
  }
  if (e != cudaSuccess) {
    if (c->base[rank]) cudaFree(c->base[rank]);
    if (c->host_status) cudaFreeHost(c->host_status);
    delete c;
    return fail(B2_ERR_CUDA, "comm_create: %s", cudaGetErrorString(e));
  }
  c->opened = (world == 1);
  *out = c;
  return B2_OK;
}

extern "C" int b2_comm_get_handle(b2_comm* c, void* handle_out) {
  using namespace b2;
  B2_REQUIRE(c && handle_out, "comm_get_handle: null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == B2_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  B2_CUDA_CHECK(cudaIpcGetMemHandle(&h, c->base[c->rank]));
  std::memcpy(handle_out, &h, sizeof(h));
  return B2_OK;
}

extern "C" int b2_comm_open_peers(b2_comm* c, const void* handles) {
  using namespace b2;
  B2_REQUIRE(c && handles, "comm_open_peers: null pointer");
  if (c->opened) return B2_OK;
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, (const char*)handles + (size_t)r * B2_IPC_HANDLE_BYTES, sizeof(h));
    B2_CUDA_CHECK(cudaIpcOpenMemHandle((void**)&c->base[r], h, cudaIpcMemLazyEnablePeerAccess));
  }
  c->opened = true;
  return B2_OK;
}

extern "C" int64_t b2_comm_max_numel(b2_comm* c, int wire_dtype) {
  if (!c) return b2::fail(B2_ERR_INVALID, "comm_max_numel: null comm");
  const size_t elt = wire_dtype == B2_BF16 ? 2 : 4;
  return (int64_t)((c->half_bytes - 64) / elt);
}

extern "C" int b2_allreduce(b2_comm* c, void* buf, int64_t numel, int wire_dtype, float scale,
                            int algo, b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(c, "allreduce: null comm");
  if (!c->opened) return fail(B2_ERR_STATE, "allreduce: b2_comm_open_peers has not been called");
  B2_REQUIRE(wire_dtype == B2_F32 || wire_dtype == B2_BF16, "allreduce: wire dtype %d", wire_dtype);
  B2_REQUIRE(algo >= B2_ALGO_AUTO && algo <= B2_ALGO_TWO_SHOT, "allreduce: algo %d", algo);
  B2_REQUIRE(numel >= 0, "allreduce: negative numel");
  if (numel == 0) return B2_OK;
  B2_REQUIRE(buf, "allreduce: null buffer");
  B2_REQUIRE((((uintptr_t)buf) & 15) == 0, "allreduce: buffer must be 16-byte aligned");
  const bool bf16w = wire_dtype == B2_BF16;
  const int epp = bf16w ? 8 : 4;
  // elements per launch: whole packs, so that every chunk start stays 16-byte aligned
  const long long cap = (b2_comm_max_numel(c, wire_dtype) / epp) * epp;
  for (long long off = 0; off < numel; off += cap) {
    const long long n = (numel - off < cap) ? (numel - off) : cap;
    const int a = algo == B2_ALGO_AUTO ? pick_algo(c->world, n * 4) : algo;
    ArArgs<1> args;
    RankView& v = args.v[0];
    fill_view_common(c, v);
    v.buf = (float*)buf + off;
    for (int r = 0; r < B2_MAX_WORLD; ++r)
      v.stage[r] = c->base[r < c->world ? r : c->rank] + kPadBytes + (size_t)c->parity * c->half_bytes;
    if (c->sumsq) {
      v.sumsq = (a == B2_ALGO_ONE_SHOT || c->world == 1) ? 1 : c->world;
      c->sumsq_numel += n;
    }
    KernelFn fn = pick_kernel<1>(c->world, bf16w, a);
    if (!fn) return fail(B2_ERR_INVALID, "allreduce: no kernel for world %d", c->world);
    const int nb = pick_blocks(n / epp, c->max_blocks, c->ppt);
    long long n_arg = n;
    float scale_arg = scale;
    uint32_t epoch = c->epoch;
    unsigned long long to = c->timeout_ns;
    void* kargs[] = {&args, &n_arg, &scale_arg, &epoch, &to};
    B2_CUDA_CHECK(cudaLaunchKernel(fn, dim3(nb), dim3(kArThreads), kargs, 0, (cudaStream_t)stream));
    c->epoch += 2;
    c->parity ^= 1;
  }
  return B2_OK;
}


// Base address and size of the cudaMalloc allocation that contains ptr (driver entry point
// resolved through the runtime, so the library has no link-time dependency on libcuda).
static int allocation_range(const void* ptr, char** base, size_t* size) {
  using namespace b2;
  typedef int (*GetRange)(unsigned long long*, size_t*, unsigned long long);
  static GetRange fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    B2_CUDA_CHECK(cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q));
    if (!f || q != cudaDriverEntryPointSuccess)
      return fail(B2_ERR_CUDA, "comm_export: cuMemGetAddressRange is not available from this driver");
    fn = (GetRange)f;
  }
  unsigned long long b = 0;
  size_t sz = 0;
  const int rc = fn(&b, &sz, (unsigned long long)(uintptr_t)ptr);
  if (rc != 0) return fail(B2_ERR_CUDA, "comm_export: cuMemGetAddressRange failed with CUresult %d", rc);
  *base = (char*)(uintptr_t)b;
  *size = sz;
  return B2_OK;
}

extern "C" int b2_comm_export(b2_comm* c, const void* ptr, void* handle_out, uint64_t* offset_out) {
  using namespace b2;
  B2_REQUIRE(c && ptr && handle_out && offset_out, "comm_export: null pointer");
  char* base = nullptr;
  size_t size = 0;
  int rc = allocation_range(ptr, &base, &size);
  if (rc != B2_OK) return rc;
  cudaIpcMemHandle_t h;
  B2_CUDA_CHECK(cudaIpcGetMemHandle(&h, base));
  std::memcpy(handle_out, &h, sizeof(h));
  *offset_out = (uint64_t)((const char*)ptr - base);
  return B2_OK;
}

extern "C" int b2_comm_register(b2_comm* c, void* local_ptr, size_t bytes, const void* handles,
                                const uint64_t* offsets, int* reg_id_out) {
  using namespace b2;
  B2_REQUIRE(c && local_ptr && reg_id_out, "comm_register: null pointer");
  B2_REQUIRE(c->world == 1 || (handles && offsets), "comm_register: null handles/offsets");
  B2_REQUIRE((((uintptr_t)local_ptr) & 15) == 0, "comm_register: buffer must be 16-byte aligned");
  if (!c->opened && c->world > 1) return fail(B2_ERR_STATE, "comm_register: peers are not opened yet");
  b2_registration reg;
  reg.live = true;
  reg.bytes = bytes;
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) {
      reg.ptr[r] = (char*)local_ptr;
      continue;
    }
    const unsigned char* h = (const unsigned char*)handles + (size_t)r * B2_IPC_HANDLE_BYTES;
    char* base = nullptr;
    for (auto& o : c->peer_allocs)
      if (o.peer == r && std::memcmp(o.handle, h, B2_IPC_HANDLE_BYTES) == 0) base = o.base;
    if (!base) {
      cudaIpcMemHandle_t ih;
      std::memcpy(&ih, h, sizeof(ih));
      B2_CUDA_CHECK(cudaIpcOpenMemHandle((void**)&base, ih, cudaIpcMemLazyEnablePeerAccess));
      b2_opened_handle o;
      o.peer = r;
      std::memcpy(o.handle, h, B2_IPC_HANDLE_BYTES);
      o.base = base;
      c->peer_allocs.push_back(o);
    }
    reg.ptr[r] = base + offsets[r];
    B2_REQUIRE((((uintptr_t)reg.ptr[r]) & 15) == 0, "comm_register: peer %d buffer is not 16-byte aligned", r);
  }
  for (size_t i = 0; i < c->regs.size(); ++i) {
    if (!c->regs[i].live) {
      c->regs[i] = reg;
      *reg_id_out = (int)i;
      return B2_OK;
    }
  }
  c->regs.push_back(reg);
  *reg_id_out = (int)c->regs.size() - 1;
  return B2_OK;
}

extern "C" int b2_comm_unregister(b2_comm* c, int reg_id) {
  using namespace b2;
  B2_REQUIRE(c && reg_id >= 0 && reg_id < (int)c->regs.size() && c->regs[reg_id].live,
             "comm_unregister: unknown registration %d", reg_id);
  c->regs[reg_id].live = false;  // mappings stay cached: the allocation may hold other buffers
  return B2_OK;
}

extern "C" int b2_allreduce_registered(b2_comm* c, int reg_id, int64_t numel, float scale,
                                       b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(c && reg_id >= 0 && reg_id < (int)c->regs.size() && c->regs[reg_id].live,
             "allreduce_registered: unknown registration %d", reg_id);
  const b2_registration& reg = c->regs[reg_id];
  B2_REQUIRE(numel >= 0 && (size_t)numel * 4 <= reg.bytes,
             "allreduce_registered: %lld elements exceed the registered %zu bytes", (long long)numel, reg.bytes);
  if (numel == 0) return B2_OK;
  ArArgs<1> args;
  RankView& v = args.v[0];
  fill_view_common(c, v);
  v.buf = (float*)reg.ptr[c->rank];
  for (int r = 0; r < B2_MAX_WORLD; ++r) v.stage[r] = reg.ptr[r < c->world ? r : c->rank];
  if (c->sumsq) {
    v.sumsq = c->world;
    c->sumsq_numel += numel;
  }
  static const bool pull = env_ll("B2_AR_ZERO_COPY_PULL", 0) != 0;  // A/B knob: two-shot pull variant
  KernelFn fn = pull ? pick_inplace_kernel<1>(c->world) : pick_push_kernel<1>(c->world);
  if (!fn) return fail(B2_ERR_INVALID, "allreduce_registered: no kernel for world %d", c->world);
  const int nb = pick_blocks(numel / 4 / c->world, c->max_blocks, c->ppt);
  long long n_arg = numel;
  float scale_arg = scale;
  uint32_t epoch = c->epoch;
  unsigned long long to = c->timeout_ns;
  int gated = c->world > 1 ? 1 : 0;
  if (gated) {
    KernelFn gate = c->world == 2 ? (KernelFn)gate_kernel<2>
                  : c->world == 4 ? (KernelFn)gate_kernel<4> : (KernelFn)gate_kernel<8>;
    void* gargs[] = {&args, &epoch, &to};
    B2_CUDA_CHECK(cudaLaunchKernel(gate, dim3(1), dim3(32), gargs, 0, (cudaStream_t)stream));
  }
  int unroll = c->unroll;
  void* kargs[] = {&args, &n_arg, &scale_arg, &epoch, &to, &gated, &unroll};  // the pull variant ignores the last
  B2_CUDA_CHECK(cudaLaunchKernel(fn, dim3(nb), dim3(kArThreads), kargs, 0, (cudaStream_t)stream));
  c->epoch += 3;  // gate, (barrier B of the pull variant,) final barrier
  return B2_OK;
}

// Whether the current device can join an NVLink multicast object (NVLS: multimem.ld_reduce /
// multimem.st). Information only in this round: the in-switch reduction path is the next step.
extern "C" int b2_nvls_supported(int* supported_out) {
  using namespace b2;
  B2_REQUIRE(supported_out, "nvls_supported: null pointer");
  *supported_out = 0;
  int dev = 0;
  B2_CUDA_CHECK(cudaGetDevice(&dev));
  typedef int (*GetAttr)(int*, int, int);
  void* f = nullptr;
  cudaDriverEntryPointQueryResult q;
  B2_CUDA_CHECK(cudaGetDriverEntryPoint("cuDeviceGetAttribute", &f, cudaEnableDefault, &q));
  if (!f || q != cudaDriverEntryPointSuccess)
    return fail(B2_ERR_CUDA, "nvls_supported: cuDeviceGetAttribute is not available");
  int v = 0;
  const int kMulticastSupported = 132;  // CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED
  const int rc = ((GetAttr)f)(&v, kMulticastSupported, dev);
  if (rc != 0) return fail(B2_ERR_CUDA, "nvls_supported: cuDeviceGetAttribute failed with CUresult %d", rc);
  *supported_out = v;
  return B2_OK;
}

static int timeout_error(const b2_comm* c) {
  return b2::fail(B2_ERR_TIMEOUT,
                  "all-reduce: a rank did not reach a cross-GPU barrier within %llu ms (B2_AR_TIMEOUT_MS); every "
                  "bucket reduced since then was filled with NaN on every rank and this communicator is unusable",
                  c->timeout_ns / 1000000ull);
}

extern "C" int b2_comm_status(b2_comm* c) {
  using namespace b2;
  B2_REQUIRE(c, "comm_status: null comm");
  uint32_t s = 0;
  B2_CUDA_CHECK(cudaMemcpy(&s, c->base[c->rank] + kStatusOff, 4, cudaMemcpyDeviceToHost));
  if (s != 0 || *(volatile uint32_t*)c->host_status != 0) return timeout_error(c);
  return B2_OK;
}

// Launch-shape knobs of the zero-copy kernels, settable at run time so that one process can sweep
// them (benchmarks/microbench.py allreduce_tune); negative values leave a knob unchanged. All
// ranks must make the same call at the same point of their call sequence.
extern "C" int b2_comm_tune(b2_comm* c, int max_blocks, int packs_per_thread, int unroll) {
  using namespace b2;
  B2_REQUIRE(c, "comm_tune: null comm");
  if (max_blocks > 0) c->max_blocks = max_blocks > kArMaxBlocks ? kArMaxBlocks : max_blocks;
  if (packs_per_thread > 0) c->ppt = packs_per_thread;
  if (unroll >= 0) c->unroll = unroll;
  return B2_OK;
}

extern "C" int b2_comm_poll(b2_comm* c) {
  using namespace b2;
  B2_REQUIRE(c, "comm_poll: null comm");
  if (*(volatile uint32_t*)c->host_status != 0) return timeout_error(c);
  return B2_OK;
}

extern "C" const void* b2_comm_status_word(b2_comm* c) { return c ? (const void*)c->host_status : nullptr; }

namespace b2 {
__global__ void sumsq_clip_kernel(double* acc, float max_norm, float* norm_out, float* coef_out) {
  const float norm = (float)sqrt(*acc);
  norm_out[0] = norm;
  const float coef = max_norm / (norm + 1e-6f);  // torch: clamp(max_norm / (total_norm + 1e-6), max=1)
  coef_out[0] = coef > 1.0f ? 1.0f : coef;
  *acc = 0.0;
}
}  // namespace b2

extern "C" int b2_comm_sumsq_enable(b2_comm* c, int on) {
  using namespace b2;
  B2_REQUIRE(c, "comm_sumsq_enable: null comm");
  c->sumsq = on != 0;
  c->sumsq_numel = 0;
  B2_CUDA_CHECK(cudaMemset(c->base[c->rank] + kAccOff, 0, 8));
  return B2_OK;
}

extern "C" int64_t b2_comm_sumsq_numel(b2_comm* c) { return c ? (int64_t)c->sumsq_numel : -1; }

extern "C" int b2_comm_sumsq_clip(b2_comm* c, float max_norm, float* norm_out, float* coef_out,
                                  b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(c && norm_out && coef_out, "comm_sumsq_clip: null pointer");
  B2_REQUIRE(max_norm > 0.f, "comm_sumsq_clip: max_norm must be positive");
  if (!c->sumsq) return fail(B2_ERR_STATE, "comm_sumsq_clip: b2_comm_sumsq_enable has not been called");
  sumsq_clip_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((double*)(c->base[c->rank] + kAccOff), max_norm, norm_out,
                                                       coef_out);
  B2_CUDA_CHECK(cudaGetLastError());
  c->sumsq_numel = 0;
  return B2_OK;
}

extern "C" int b2_comm_destroy(b2_comm* c) {
  using namespace b2;
  if (!c) return B2_OK;
  b2_nvls_release(c);
  for (auto& h : c->peer_allocs) cudaIpcCloseMemHandle(h.base);
  for (int r = 0; r < c->world; ++r) {
    if (!c->base[r]) continue;
    if (r == c->rank) cudaFree(c->base[r]); else cudaIpcCloseMemHandle(c->base[r]);
  }
  if (c->host_status) cudaFreeHost(c->host_status);
  delete c;
  return B2_OK;
}

extern "C" size_t b2_allreduce_emulated_scratch_bytes(int world, int64_t numel) {
  if (world < 1 || numel < 0) return 0;
  const size_t half = (((size_t)numel * 4 + 64) + 255) & ~(size_t)255;
  return (size_t)world * (b2::kPadBytes + half);
}

extern "C" int b2_allreduce_emulated(void* const* bufs_host, int world, int64_t numel,
                                     int wire_dtype, float scale, int algo, void* staging,
                                     b2_stream_t stream) {
  return b2_allreduce_emulated_sumsq(bufs_host, world, numel, wire_dtype, scale, algo, staging, nullptr, stream);
}

extern "C" int b2_allreduce_emulated_sumsq(void* const* bufs_host, int world, int64_t numel,
                                           int wire_dtype, float scale, int algo, void* staging,
                                           double* sumsq_out, b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(bufs_host && staging, "allreduce_emulated: null pointer");
  B2_REQUIRE(world == 1 || world == 2 || world == 4 || world == 8, "allreduce_emulated: world %d", world);
  B2_REQUIRE(algo >= B2_ALGO_ONE_SHOT && algo <= B2_ALGO_PUSH_INPLACE, "allreduce_emulated: algo %d", algo);
  const bool inplace = (algo == B2_ALGO_TWO_SHOT_INPLACE || algo == B2_ALGO_PUSH_INPLACE);
  B2_REQUIRE(!inplace || wire_dtype == B2_F32,
             "allreduce_emulated: the in-place algorithms are fp32 on the wire");
  B2_REQUIRE(wire_dtype == B2_F32 || wire_dtype == B2_BF16, "allreduce_emulated: wire dtype");
  B2_REQUIRE(numel > 0, "allreduce_emulated: numel must be positive");
  const bool bf16w = wire_dtype == B2_BF16;
  const int epp = bf16w ? 8 : 4;
  const size_t half = (((size_t)numel * 4 + 64) + 255) & ~(size_t)255;
  const size_t per_rank = kPadBytes + half;
  cudaStream_t st = (cudaStream_t)stream;
  for (int r = 0; r < world; ++r)
    B2_CUDA_CHECK(cudaMemsetAsync((char*)staging + r * per_rank, 0, kPadBytes, st));
  ArArgs<B2_MAX_WORLD> args;
  for (int r = 0; r < B2_MAX_WORLD; ++r) {
    RankView& v = args.v[r];
    const int me = r < world ? r : 0;
    B2_REQUIRE(bufs_host[me] && (((uintptr_t)bufs_host[me]) & 15) == 0,
               "allreduce_emulated: buffer %d null or not 16-byte aligned", me);
    v.buf = (float*)bufs_host[me];
    v.rank = me;
    for (int k = 0; k < B2_MAX_WORLD; ++k) {
      char* base = (char*)staging + (size_t)(k < world ? k : me) * per_rank;
      v.stage[k] = inplace ? (char*)bufs_host[k < world ? k : me] : base + kPadBytes;
      v.flags[k] = (uint32_t*)base;
    }
    v.status = (uint32_t*)((char*)staging + (size_t)me * per_rank + kStatusOff);
    v.host_status = v.status;  // no host mirror in the emulation
    v.sumsq = !sumsq_out ? 0 : (algo == B2_ALGO_ONE_SHOT || world == 1) ? 1 : world;
  }
  KernelFn fn = (algo == B2_ALGO_TWO_SHOT_INPLACE) ? pick_inplace_kernel<B2_MAX_WORLD>(world)
              : (algo == B2_ALGO_PUSH_INPLACE)     ? pick_push_kernel<B2_MAX_WORLD>(world)
                                                   : pick_kernel<B2_MAX_WORLD>(world, bf16w, algo);
  // The virtual ranks' CTAs wait on one another, so they must all be resident: cooperative
  // launch, and few enough CTAs (<= 1 per SM with 512 threads is always resident on 148 SMs).
  int nb = pick_blocks(numel / epp, 16);
  long long n_arg = numel;
  float scale_arg = scale;
  uint32_t epoch = 1;
  unsigned long long to = 5ull * 1000000000ull;
  int gated = 0;
  int unroll = 0;
  void* kargs[] = {&args, &n_arg, &scale_arg, &epoch, &to, &gated, &unroll};  // staging kernels ignore the extra slots
  B2_CUDA_CHECK(cudaLaunchCooperativeKernel(fn, dim3(nb, world), dim3(kArThreads), kargs, 0, st));
  if (sumsq_out)
    for (int r = 0; r < world; ++r)
      B2_CUDA_CHECK(cudaMemcpyAsync(sumsq_out + r, (char*)staging + (size_t)r * per_rank + kAccOff, 8,
                                    cudaMemcpyDeviceToDevice, st));
  return B2_OK;
}
