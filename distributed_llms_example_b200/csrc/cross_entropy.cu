// Label-masked cross-entropy forward / backward (include/b2ddp.h, section (c)).
//
// Replaces CrossEntropyLoss(ignore_index=-100) inside T5 (SP/transformers/models/t5/
// modeling_t5.py:1112-1117; reached from train-accelerator.py:220-222, train-task.py:289-292,
// train-torchrun.py:140), i.e. ATen's _log_softmax + nll_loss_forward / nll_loss_backward +
// _log_softmax_backward_data and, under bf16 autocast, the fp32 up-cast copy before them.
//
// Forward  : one pass over the logits (N*V*s bytes). The unit of work is a WARP: it streams a
//            row (or 1/2..1/16 of a row while N is too small to keep 148 SMs busy) with
//            coalesced 16-byte loads, keeps an online (max, sum-exp) per lane in registers,
//            merges with warp shuffles only - no CTA barrier on the data path - and the last
//            CTA to arrive sums the row losses in a fixed order (scratch: arrival counter, row
//            losses, split partials). Shared memory only stages the final block reduction.
//            (A CTA-per-row version with a shared-memory combine reached 70 % of the HBM copy
//            peak for bf16; ncu showed the CTA barrier stalling as many issue slots as memory;
//            this version reaches 92 %.)
// Backward : one read + one write of N*V elements, CTA per (row, 1024-pack tile); ignored rows
//            are written as zeros without reading their logits.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace b2 {

constexpr int kCeThreads = 256;
constexpr int kCeMaxSplit = 16;

// (max, sum) of exp(x - max), the running state of an online log-sum-exp.
struct __align__(8) MS {
  float m, s;
};

__device__ __forceinline__ float safe_ref(float m) { return m == -INFINITY ? 0.f : m; }

__device__ __forceinline__ MS ms_merge(MS a, MS b) {
  const float m = fmaxf(a.m, b.m);
  const float r = safe_ref(m);
  MS o;
  o.m = m;
  o.s = a.s * fast_exp2((a.m - r) * kLog2e) + b.s * fast_exp2((b.m - r) * kLog2e);
  return o;
}

template <int NV>
__device__ __forceinline__ void ms_update(MS& st, const float (&x)[NV]) {
  float lm = x[0];
#pragma unroll
  for (int i = 1; i < NV; ++i) lm = fmaxf(lm, x[i]);
  const float m = fmaxf(st.m, lm);
  const float r = safe_ref(m);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) acc += fast_exp2((x[i] - r) * kLog2e);
  st.s = st.s * fast_exp2((st.m - r) * kLog2e) + acc;
  st.m = m;
}

template <bool BF16>
struct Elem;
template <>
struct Elem<true> {
  using T = __nv_bfloat16;
  static constexpr int kPerPack = 8;
  static __device__ __forceinline__ void unpack(uint4 u, float* x) {
    x[0] = bf16lo(u.x); x[1] = bf16hi(u.x); x[2] = bf16lo(u.y); x[3] = bf16hi(u.y);
    x[4] = bf16lo(u.z); x[5] = bf16hi(u.z); x[6] = bf16lo(u.w); x[7] = bf16hi(u.w);
  }
  static __device__ __forceinline__ uint4 pack(const float* x) {
    return make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]),
                      pack_bf16(x[6], x[7]));
  }
  static __device__ __forceinline__ float load1(const T* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void store1(T* p, float v) { *p = __float2bfloat16_rn(v); }
};
template <>
struct Elem<false> {
  using T = float;
  static constexpr int kPerPack = 4;
  static __device__ __forceinline__ void unpack(uint4 u, float* x) {
    x[0] = __uint_as_float(u.x); x[1] = __uint_as_float(u.y);
    x[2] = __uint_as_float(u.z); x[3] = __uint_as_float(u.w);
  }
  static __device__ __forceinline__ uint4 pack(const float* x) {
    return make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]), __float_as_uint(x[2]),
                      __float_as_uint(x[3]));
  }
  static __device__ __forceinline__ float load1(const T* p) { return *p; }
  static __device__ __forceinline__ void store1(T* p, float v) { *p = v; }
};

// Scratch layout (b2_ce_workspace_bytes): [0,16) arrival counter; row_loss[N]; partial[N*split].
struct CeScratch {
  unsigned int* counter;
  float* row_loss;
  MS* partial;
};
__host__ __device__ inline CeScratch carve(void* ws, long long n_rows) {
  CeScratch s;
  char* b = (char*)ws;
  s.counter = (unsigned int*)b;
  s.row_loss = (float*)(b + 16);
  const size_t off = 16 + (((size_t)n_rows * 4 + 15) & ~(size_t)15);
  s.partial = (MS*)(b + off);
  return s;
}

__device__ __forceinline__ float block_sum(float v, float* smem) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    constexpr int kWarps = kCeThreads / 32;
    v = lane < kWarps ? smem[lane] : 0.f;
    v = warp_sum(v);
  }
  return v;  // valid in thread 0
}

// Largest element of U raw 16-byte packs. bf16: packed HMNMX2 on the raw words (max is exact in
// bf16), one unpack at the end; fp32: plain fmaxf chain (ptxas fuses pairs into 3-input FMNMX).
template <bool BF16, int U>
__device__ __forceinline__ float packs_max(const uint4 (&r)[U]) {
  if constexpr (BF16) {
    __nv_bfloat162 m = *reinterpret_cast<const __nv_bfloat162*>(&r[0].x);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t w[4] = {r[u].x, r[u].y, r[u].z, r[u].w};
#pragma unroll
      for (int k = 0; k < 4; ++k) m = __hmax2(m, *reinterpret_cast<const __nv_bfloat162*>(&w[k]));
    }
    const uint32_t mw = *reinterpret_cast<uint32_t*>(&m);
    return fmaxf(bf16lo(mw), bf16hi(mw));
  } else {
    float m = __uint_as_float(r[0].x);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      m = fmaxf(m, fmaxf(__uint_as_float(r[u].x), __uint_as_float(r[u].y)));
      m = fmaxf(m, fmaxf(__uint_as_float(r[u].z), __uint_as_float(r[u].w)));
    }
    return m;
  }
}

// 2^x for x <= 0 on the FMA and integer pipes only (no MUFU): round-to-nearest split x = n + f with
// the 1.5*2^23 trick, a degree-4 polynomial for 2^f on [-0.5, 0.5] (max relative error 3.7e-6,
// far inside the 1e-5 bar of the parity tests), then n added into the exponent field. One element
// in four of the bf16 forward goes this way: with one MUFU.EX2 per element the XU pipe (16/clk/SM)
// caps the kernel near 6 TB/s of bf16 logits, below what HBM delivers (ncu: pipe_xu 72 % busy at
// 4.4 TB/s); with a quarter of the exponentials moved over, XU and FMA pipes are balanced above
// the HBM rate. x < -126 gives 0; NaN propagates like the MUFU path.
__device__ __forceinline__ float poly_exp2(float x) {
  const float kMagic = 12582912.f;  // 1.5 * 2^23
  const float t = __fadd_rn(x, kMagic);
  const float n = __fsub_rn(t, kMagic);
  const float f = __fsub_rn(x, n);
  float p = __fmaf_rn(0.009676037356257439f, f, 0.05592203512787819f);
  p = __fmaf_rn(p, f, 0.2402210682630539f);
  p = __fmaf_rn(p, f, 0.6931210160255432f);
  p = __fmaf_rn(p, f, 1.0000001192092896f);
  const float r = __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
  return x < -126.f ? 0.f : r;
}

// st <- st merged with the U packs in r: one FFMA + one MUFU.EX2 + one FADD per element.
// exp(x - ref) is evaluated as ex2(x*log2e - ref*log2e); NaN inputs propagate (hmax2/fmaxf drop
// NaN from the max, but ex2(NaN) lands in the sum).
template <bool BF16, int U, bool POLY = false>
__device__ __forceinline__ void ms_update_packs(MS& st, const uint4 (&r)[U]) {
  using E = Elem<BF16>;
  const float m = fmaxf(st.m, packs_max<BF16, U>(r));
  const float ref = safe_ref(m);
  const float neg = -ref * kLog2e;
  float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    float x[E::kPerPack];
    E::unpack(r[u], x);
#pragma unroll
    for (int k = 0; k < E::kPerPack; k += 2) {
      acc0 += fast_exp2(fmaf(x[k], kLog2e, neg));
      if (POLY && (k & 2))  // elements 3 and 7 of every pack of 8 take the FMA-pipe polynomial
        acc1 += poly_exp2(fmaf(x[k + 1], kLog2e, neg));
      else
        acc1 += fast_exp2(fmaf(x[k + 1], kLog2e, neg));
    }
  }
  st.s = st.s * fast_exp2(fmaf(st.m, kLog2e, neg)) + (acc0 + acc1);
  st.m = m;
}

constexpr int kCeFwdCtasPerSM = 4;
constexpr int kCeWarps = kCeThreads / 32;

__device__ __forceinline__ MS warp_merge(MS st) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MS other;
    other.m = __shfl_xor_sync(0xffffffffu, st.m, o);
    other.s = __shfl_xor_sync(0xffffffffu, st.s, o);
    st = ms_merge(st, other);
  }
  return st;
}

// The tail of ce_forward_kernel, run by the last CTA to arrive. Kept out of line so that its
// registers (a batch of rows and all partials of a row in flight) do not raise the pressure in the
// streaming loop, which is compiled for 4 CTAs per SM.
__device__ __noinline__ void ce_fold_rows(const CeScratch sc, const long long* __restrict__ labels,
                                          long long n_rows, int vocab, long long ignore_index, int split,
                                          float* __restrict__ row_lse, float* acc_out, float* cnt_out) {
  // Last CTA to arrive: (split rows only) fold the partials of each row, then sum the valid
  // rows' losses in a fixed order (thread-strided, then a tree), so the result does not
  // depend on which CTA happened to be last. Everything read here is coalesced, and the loads
  // of kFoldRows rows (and of all partials of a row) are issued before the first one is used:
  // written as a plain loop this tail cost one L2 round trip per row per thread, ~20-30 us of a
  // 35-120 us kernel (N <= 8192); batched it is a handful of round trips.
  constexpr int kFoldRows = 8;
  float acc = 0.f, cnt = 0.f;
  for (long long r0 = threadIdx.x; r0 < n_rows; r0 += (long long)kCeThreads * kFoldRows) {
    long long ys[kFoldRows];
    float ls[kFoldRows];
#pragma unroll
    for (int i = 0; i < kFoldRows; ++i) {
      const long long r = r0 + (long long)i * kCeThreads;
      ys[i] = ignore_index;
      ls[i] = 0.f;
      if (r < n_rows) {
        ys[i] = labels[r];
        ls[i] = __ldcg(sc.row_loss + r);
      }
    }
    if (split != 1) {
#pragma unroll 2
      for (int i = 0; i < kFoldRows; ++i) {
        const long long r = r0 + (long long)i * kCeThreads;
        if (r >= n_rows) break;
        float2 q[kCeMaxSplit];
#pragma unroll
        for (int k = 0; k < kCeMaxSplit; ++k)
          if (k < split) q[k] = __ldcg(reinterpret_cast<const float2*>(sc.partial + r * split + k));
        MS t{-INFINITY, 0.f};
#pragma unroll
        for (int k = 0; k < kCeMaxSplit; ++k)
          if (k < split) t = ms_merge(t, MS{q[k].x, q[k].y});
        const float lse = t.m + logf(t.s);
        row_lse[r] = lse;
        ls[i] = (ys[i] < 0 || ys[i] >= vocab) ? __int_as_float(0x7fc00000) : lse - ls[i];
      }
    }
#pragma unroll
    for (int i = 0; i < kFoldRows; ++i) {  // same rows, same order as a plain strided loop
      if (ys[i] != ignore_index) {
        acc += ls[i];
        cnt += 1.f;
      }
    }
  }
  *acc_out = acc;
  *cnt_out = cnt;
}

// The unit of work is a WARP, not a CTA: warp w of the grid walks work items (row, part) with
// stride (gridDim.x * 8). A warp streams its piece with 512-byte coalesced requests, merges with
// shuffles only and never waits for another warp, so there is no CTA barrier on the data path
// (an earlier CTA-per-row version spent as many issue slots stalled on bar.sync as on memory).
// Loads for the next batch of packs are issued before the current batch is reduced.
template <bool BF16, bool POLY>
__global__ void __launch_bounds__(kCeThreads, kCeFwdCtasPerSM)
ce_forward_kernel(const typename Elem<BF16>::T* __restrict__ logits,
                  const long long* __restrict__ labels, long long n_rows, int vocab,
                  long long row_stride, long long ignore_index, int split,
                  float* __restrict__ row_lse, float* __restrict__ loss_out,
                  float* __restrict__ nvalid_out, void* __restrict__ ws) {
  using E = Elem<BF16>;
  using T = typename E::T;
  constexpr int PP = E::kPerPack;
  constexpr int U = 4;
  __shared__ float s_f[kCeWarps];
  __shared__ bool s_last;

  const CeScratch sc = carve(ws, n_rows);
  const long long items = n_rows * split;
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * kCeWarps + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * kCeWarps;
  for (long long item = warp0; item < items; item += nwarps) {
    const long long row = item / split;
    const int part = (int)(item - row * split);
    const T* zrow = logits + row * row_stride;

    // Lane 0 fetches the label and the label's logit now; they are needed only after the
    // piece has been streamed, so their latency hides behind the main loop.
    long long y = ignore_index;
    float zy = 0.f;
    if (lane == 0 && part == 0) {
      y = labels[row];
      if (y != ignore_index && y >= 0 && y < vocab) zy = E::load1(zrow + y);
    }

    const bool vec_ok = ((((uintptr_t)zrow) & 15) == 0);
    const int npack = vec_ok ? vocab / PP : 0;
    const int per = (npack + split - 1) / split;
    const int p0 = min(part * per, npack), p1 = min(p0 + per, npack);
    const uint4* zp = reinterpret_cast<const uint4*>(zrow);

    MS st{-INFINITY, 0.f};
    int i = p0 + lane;
    constexpr int STEP = U * 32;
    if (i + (U - 1) * 32 < p1) {
      uint4 cur[U];
#pragma unroll
      for (int u = 0; u < U; ++u) cur[u] = ld_nc16(zp + i + u * 32);
      for (i += STEP; i + (U - 1) * 32 < p1; i += STEP) {
        uint4 nxt[U];
#pragma unroll
        for (int u = 0; u < U; ++u) nxt[u] = ld_nc16(zp + i + u * 32);
        ms_update_packs<BF16, U, POLY>(st, cur);
#pragma unroll
        for (int u = 0; u < U; ++u) cur[u] = nxt[u];
      }
      ms_update_packs<BF16, U, POLY>(st, cur);
    }
    for (; i < p1; i += 32) {
      const uint4 one[1] = {ld_nc16(zp + i)};
      ms_update_packs<BF16, 1, POLY>(st, one);
    }
    // Scalar tail (vocab % PP, or the whole row if it is not 16-byte aligned): last part only.
    if (part == split - 1) {
      for (int j = npack * PP + lane; j < vocab; j += 32) {
        float x[1] = {E::load1(zrow + j)};
        ms_update<1>(st, x);
      }
    }
    st = warp_merge(st);

    if (lane == 0) {
      if (split == 1) {
        // Whole row seen by this warp: finish the row here so the last CTA only has to sum.
        const float lse = st.m + logf(st.s);
        row_lse[row] = lse;
        float l = 0.f;
        if (y != ignore_index)
          l = (y < 0 || y >= vocab) ? __int_as_float(0x7fc00000)  // out-of-range label poisons the loss
                                    : lse - zy;
        sc.row_loss[row] = l;
      } else {
        sc.partial[row * split + part] = st;
        if (part == 0) sc.row_loss[row] = zy;  // split rows: the slot carries z[y] to the fold below
      }
    }
  }
  __syncthreads();  // every warp of this CTA has written its results
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int arrived = atomicAdd(sc.counter, 1u);
    s_last = (arrived == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();

  float acc, cnt;
  ce_fold_rows(sc, labels, n_rows, vocab, ignore_index, split, row_lse, &acc, &cnt);
  acc = block_sum(acc, s_f);
  cnt = block_sum(cnt, s_f);
  if (threadIdx.x == 0) {
    loss_out[0] = acc / cnt;  // 0/0 = NaN when every row is ignored, as torch
    nvalid_out[0] = cnt;
    *sc.counter = 0u;  // ready for the next call on this stream
  }
}

constexpr int kCeBwdCtasPerSM = 6;
constexpr int kCeBwdU = 4;  // packs per thread per tile

// Persistent over (row, tile) work items; a tile is kCeBwdU * 256 packs of one row.
template <bool BF16>
__global__ void __launch_bounds__(kCeThreads, kCeBwdCtasPerSM)
ce_backward_kernel(const typename Elem<BF16>::T* __restrict__ logits,
                   const long long* __restrict__ labels, const float* __restrict__ row_lse,
                   const float* __restrict__ grad_out, const float* __restrict__ nvalid,
                   long long n_rows, int vocab, long long row_stride, long long ignore_index,
                   int tiles_per_row, typename Elem<BF16>::T* __restrict__ dlogits) {
  using E = Elem<BF16>;
  using T = typename E::T;
  constexpr int PP = E::kPerPack;
  constexpr int U = kCeBwdU;
  const float scale = __ldg(grad_out) / __ldg(nvalid);
  const long long items = n_rows * tiles_per_row;
  for (long long item = blockIdx.x; item < items; item += gridDim.x) {
    const long long row = item / tiles_per_row;
    const int tile = (int)(item - row * tiles_per_row);
    const T* zrow = logits + row * row_stride;
    T* drow = dlogits + row * row_stride;
    const long long y = __ldg(labels + row);
    const bool valid = (y != ignore_index);
    const float coef = valid ? scale : 0.f;
    const float neg = -__ldg(row_lse + row) * kLog2e;

    const bool vec_ok = (((((uintptr_t)zrow) | ((uintptr_t)drow)) & 15) == 0);
    const int npack = vec_ok ? vocab / PP : 0;
    const int per = U * kCeThreads;
    const int p0 = min(tile * per, npack), p1 = min(p0 + per, npack);
    const uint4* zp = reinterpret_cast<const uint4*>(zrow);
    uint4* dp = reinterpret_cast<uint4*>(drow);

    if (!valid) {
      const uint4 zero = make_uint4(0, 0, 0, 0);
      for (int i = p0 + threadIdx.x; i < p1; i += kCeThreads) dp[i] = zero;
    } else {
      uint4 r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = p0 + threadIdx.x + u * kCeThreads;
        if (i < p1) r[u] = ld_nc16(zp + i);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = p0 + threadIdx.x + u * kCeThreads;
        if (i < p1) {
          float x[PP];
          E::unpack(r[u], x);
#pragma unroll
          for (int k = 0; k < PP; ++k) x[k] = fast_exp2(fmaf(x[k], kLog2e, neg)) * coef;
          dp[i] = E::pack(x);
        }
      }
      // The label's column lies in exactly one pack of one thread of one tile. That thread
      // overwrites the element after its vector store (same thread, same address: program
      // order holds), which keeps the main loop free of per-column compares.
      asm volatile("" ::: "memory");  // keep the compiler from moving the patch above the stores
      const long long yp = y / PP;
      if (y >= 0 && y < (long long)npack * PP && yp >= p0 && yp < p1 &&
          (int)((yp - p0) % kCeThreads) == (int)threadIdx.x) {
        const float p = fast_exp2(fmaf(E::load1(zrow + y), kLog2e, neg));
        E::store1(drow + y, (p - 1.f) * coef);
      }
    }
    if (tile == tiles_per_row - 1) {
      for (int j = npack * PP + threadIdx.x; j < vocab; j += kCeThreads) {
        float o = 0.f;
        if (valid) {
          const float p = fast_exp2(fmaf(E::load1(zrow + j), kLog2e, neg));
          o = (p - ((long long)j == y ? 1.f : 0.f)) * coef;
        }
        E::store1(drow + j, o);
      }
    }
  }
}

// CTAs per SM the persistent grids are capped at; 0 = one CTA per work item (no cap).
// B2_CE_POLY=1 moves a quarter of the bf16 forward's exponentials from MUFU.EX2 to poly_exp2 (read
// on every call so that a benchmark can A/B the two in one process). Default off: see
// profiles/README.md for the measured A/B.
static bool ce_poly() {
  const char* s = std::getenv("B2_CE_POLY");
  return s && s[0] == '1';
}

// B2_CE_FWD_CTAS_PER_SM / B2_CE_BWD_CTAS_PER_SM override the defaults for tuning runs.
static int env_int(const char* name, int dflt) {
  const char* v = std::getenv(name);
  return (v && *v) ? std::atoi(v) : dflt;
}
static long long fwd_grid(long long items) {  // items are per-warp work units, 8 warps per CTA
  const int cap = env_int("B2_CE_FWD_CTAS_PER_SM", 0);  // read per call: benchmarks sweep it in one process
  const long long ctas = (items + kCeWarps - 1) / kCeWarps;
  return cap > 0 ? std::min<long long>(ctas, (long long)kNumSMs * cap) : ctas;
}
static long long bwd_grid(long long items) {
  static const int cap = env_int("B2_CE_BWD_CTAS_PER_SM", 0);
  return cap > 0 ? std::min<long long>(items, (long long)kNumSMs * cap) : items;
}

static int pick_split(long long n_rows) {
  // A row is split over several warps only while there are fewer rows than ~2048 warps' worth of
  // work: measured on B200 (profiles/README.md, round 2 sweep), one warp per row wins from 2048 rows
  // up - split rows pay for the partials and the last CTA's fold (N = 8192: 100.8 us unsplit, 122.2 us
  // split in 2; N = 2048: 33.7 vs 47.8 us split in 8) - and 2048 work items is where N = 1024 is
  // fastest (25.6 us split in 2, 29.7 us unsplit, 30.0 us split in 8).
  const int forced = env_int("B2_CE_SPLIT", 0);  // tuning override (power of two <= 16)
  if (forced >= 1 && forced <= kCeMaxSplit && (forced & (forced - 1)) == 0) return forced;
  int s = 1;
  while (s < kCeMaxSplit && n_rows * s < 2048) s <<= 1;
  return s;
}

}  // namespace b2

extern "C" size_t b2_ce_workspace_bytes(int64_t n_rows, int64_t vocab) {
  (void)vocab;
  if (n_rows < 0) n_rows = 0;
  return 16 + (((size_t)n_rows * 4 + 15) & ~(size_t)15) +
         (size_t)n_rows * b2::kCeMaxSplit * sizeof(b2::MS);
}

extern "C" int b2_ce_forward(const void* logits, const int64_t* labels, int64_t n_rows,
                             int64_t vocab, int64_t row_stride, int64_t ignore_index, int dtype,
                             float* row_lse, float* loss_out, float* nvalid_out, void* workspace,
                             size_t workspace_bytes, b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "ce_forward: dtype %d not f32/bf16", dtype);
  B2_REQUIRE(n_rows >= 0 && vocab > 0 && vocab < (1ll << 30), "ce_forward: bad shape [%lld, %lld]",
             (long long)n_rows, (long long)vocab);
  B2_REQUIRE(row_stride >= vocab, "ce_forward: row_stride %lld < vocab %lld", (long long)row_stride,
             (long long)vocab);
  B2_REQUIRE(loss_out && nvalid_out, "ce_forward: null output");
  if (n_rows == 0) {
    // mean over zero rows: NaN, n_valid 0 (torch returns nan for an empty batch too)
    static const float kNan = __builtin_nanf(""), kZero = 0.f;
    B2_CUDA_CHECK(cudaMemcpyAsync(loss_out, &kNan, 4, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    B2_CUDA_CHECK(cudaMemcpyAsync(nvalid_out, &kZero, 4, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return B2_OK;
  }
  B2_REQUIRE(logits && labels && row_lse && workspace, "ce_forward: null pointer");
  B2_REQUIRE(n_rows * kCeMaxSplit < (1ll << 31), "ce_forward: too many rows");
  B2_REQUIRE(workspace_bytes >= b2_ce_workspace_bytes(n_rows, vocab),
             "ce_forward: workspace %zu B < required %zu B", workspace_bytes,
             b2_ce_workspace_bytes(n_rows, vocab));
  B2_REQUIRE((((uintptr_t)workspace) & 15) == 0, "ce_forward: workspace must be 16-byte aligned");
  const int split = pick_split(n_rows);
  const long long grid = fwd_grid(n_rows * split);
  if (dtype == B2_BF16 && ce_poly()) {
    ce_forward_kernel<true, true><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, workspace);
  } else if (dtype == B2_BF16) {
    ce_forward_kernel<true, false><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, workspace);
  } else {
    ce_forward_kernel<false, false><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const float*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, workspace);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_ce_backward(const void* logits, const int64_t* labels, const float* row_lse,
                              const float* grad_out, const float* nvalid, int64_t n_rows,
                              int64_t vocab, int64_t row_stride, int64_t ignore_index, int dtype,
                              void* dlogits, b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "ce_backward: dtype %d not f32/bf16", dtype);
  B2_REQUIRE(n_rows >= 0 && vocab > 0 && vocab < (1ll << 30), "ce_backward: bad shape [%lld, %lld]",
             (long long)n_rows, (long long)vocab);
  B2_REQUIRE(row_stride >= vocab, "ce_backward: row_stride < vocab");
  if (n_rows == 0) return B2_OK;
  B2_REQUIRE(logits && labels && row_lse && grad_out && nvalid && dlogits, "ce_backward: null pointer");
  const int pp = dtype == B2_BF16 ? 8 : 4;
  const long long npack = vocab / pp;
  const int per = kCeBwdU * kCeThreads;
  long long tiles = (npack + per - 1) / per;
  if (tiles < 1) tiles = 1;
  B2_REQUIRE(n_rows * tiles < (1ll << 31), "ce_backward: too many tiles");
  const long long grid = bwd_grid(n_rows * tiles);
  if (dtype == B2_BF16) {
    ce_backward_kernel<true><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, row_lse, grad_out, nvalid, n_rows,
        (int)vocab, row_stride, ignore_index, (int)tiles, (__nv_bfloat16*)dlogits);
  } else {
    ce_backward_kernel<false><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const float*)logits, (const long long*)labels, row_lse, grad_out, nvalid, n_rows,
        (int)vocab, row_stride, ignore_index, (int)tiles, (float*)dlogits);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}

// ---------------------------------------------------------------------------------------
// b2_ce_forward_noalloc: the forward above for hosts that must not allocate.
//
// Some programs that embed this library cannot give it a workspace: a serving or training
// runtime that reserves the whole device for its own pool at start-up (a later cudaMalloc
// fails, or is forbidden by policy once the job is "warm"), or a thin foreign-language
// binding that has no place to keep a buffer alive between calls. For them the scratch
// lives inside the library image itself: a file-scope __device__ array of fixed size,
// present on every GPU that loads libb2ddp.so until the process exits.
//
// The array's name is only a host-side handle for the symbol, not a device address, so the
// host asks the runtime for the address once per device (cudaGetSymbolAddress, with that
// device current) and hands THAT pointer to the kernel. The first call on a device performs
// this set-up; make that first call before any stream capture begins.
//
// All calls on a device share the one array and it keeps whatever the previous call left
// in it, so the arrival counter is cleared on the stream before every launch. Two calls
// whose kernels overlap on the GPU (different streams, no ordering between them) race on
// the scratch and produce garbage: serialise them. No cudaMalloc / cudaFree / allocating
// library routine is called on this path.
// ---------------------------------------------------------------------------------------

namespace b2 {

constexpr size_t kNoAllocScratchBytes =
    16 + (size_t)B2_CE_NOALLOC_MAX_ROWS * 4 +
    (size_t)B2_CE_NOALLOC_MAX_ROWS * kCeMaxSplit * sizeof(MS);  // 1 081 360 B

static __device__ __align__(256) unsigned char g_ce_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_noalloc_ptr[kMaxDevices] = {nullptr};  // device address of the array, per device

// Device address of the scratch array on the current device (resolved once per device).
static int noalloc_scratch(void** out) {
  int dev = -1;
  B2_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(B2_ERR_INVALID, "ce_forward_noalloc: device ordinal %d outside [0, %d)", dev, kMaxDevices);
  void* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    B2_CUDA_CHECK(cudaGetSymbolAddress(&p, g_ce_noalloc_scratch));
    if (!p) return fail(B2_ERR_CUDA, "ce_forward_noalloc: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);  // same value from any thread
  }
  *out = p;
  return B2_OK;
}

}  // namespace b2

extern "C" int b2_ce_forward_noalloc(const void* logits, const int64_t* labels, int64_t n_rows,
                                     int64_t vocab, int64_t row_stride, int64_t ignore_index,
                                     int dtype, float* row_lse, float* loss_out, float* nvalid_out,
                                     b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "ce_forward_noalloc: dtype %d not f32/bf16", dtype);
  B2_REQUIRE(n_rows >= 0 && vocab > 0 && vocab < (1ll << 30), "ce_forward_noalloc: bad shape [%lld, %lld]",
             (long long)n_rows, (long long)vocab);
  if (n_rows > B2_CE_NOALLOC_MAX_ROWS)
    return fail(B2_ERR_TOO_LARGE,
                "ce_forward_noalloc: %lld rows exceed the built-in scratch (max %d rows); use "
                "b2_ce_forward with a caller-provided workspace",
                (long long)n_rows, B2_CE_NOALLOC_MAX_ROWS);
  B2_REQUIRE(row_stride >= vocab, "ce_forward_noalloc: row_stride %lld < vocab %lld",
             (long long)row_stride, (long long)vocab);
  B2_REQUIRE(loss_out && nvalid_out, "ce_forward_noalloc: null output");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_rows == 0) {
    static const float kNan = __builtin_nanf(""), kZero = 0.f;
    B2_CUDA_CHECK(cudaMemcpyAsync(loss_out, &kNan, 4, cudaMemcpyHostToDevice, st));
    B2_CUDA_CHECK(cudaMemcpyAsync(nvalid_out, &kZero, 4, cudaMemcpyHostToDevice, st));
    return B2_OK;
  }
  B2_REQUIRE(logits && labels && row_lse, "ce_forward_noalloc: null pointer");
  void* ws = nullptr;
  int rc = noalloc_scratch(&ws);
  if (rc != B2_OK) return rc;
  // The array still holds the previous call's contents; only the counter has to be zero.
  B2_CUDA_CHECK(cudaMemsetAsync(ws, 0, 16, st));
  const int split = pick_split(n_rows);
  const long long grid = fwd_grid(n_rows * split);
  if (dtype == B2_BF16 && ce_poly()) {
    ce_forward_kernel<true, true><<<(unsigned)grid, kCeThreads, 0, st>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, ws);
  } else if (dtype == B2_BF16) {
    ce_forward_kernel<true, false><<<(unsigned)grid, kCeThreads, 0, st>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, ws);
  } else {
    ce_forward_kernel<false, false><<<(unsigned)grid, kCeThreads, 0, st>>>(
        (const float*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, ws);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}
