// Label-masked cross-entropy forward / backward (include/b2ddp.h, section (c)).
//
// Replaces CrossEntropyLoss(ignore_index=-100) inside T5 (SP/transformers/models/t5/
// modeling_t5.py:1112-1117; reached from train-accelerator.py:220-222, train-task.py:289-292,
// train-torchrun.py:140), i.e. ATen's _log_softmax + nll_loss_forward / nll_loss_backward +
// _log_softmax_backward_data and, under bf16 autocast, the fp32 up-cast copy before them.
//
// Forward  : one pass over the logits (N*V*s bytes), online max/sum-exp per thread in
//            registers with 16-byte loads, warp-shuffle + shared-memory combine per CTA,
//            optional split of a row over several CTAs when N is too small to fill 148 SMs,
//            and a deterministic last-CTA reduction of the per-row losses (scratch: arrival
//            counter + row losses + split partials).
// Backward : one read + one write of N*V elements; ignored rows are written as zeros
//            without reading their logits.
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

__device__ __forceinline__ MS block_merge(MS st, MS* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MS other;
    other.m = __shfl_xor_sync(0xffffffffu, st.m, o);
    other.s = __shfl_xor_sync(0xffffffffu, st.s, o);
    st = ms_merge(st, other);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) smem[warp] = st;
  __syncthreads();
  if (warp == 0) {
    constexpr int kWarps = kCeThreads / 32;
    MS v = lane < kWarps ? smem[lane] : MS{-INFINITY, 0.f};
#pragma unroll
    for (int o = kWarps / 2; o > 0; o >>= 1) {
      MS other;
      other.m = __shfl_xor_sync(0xffffffffu, v.m, o);
      other.s = __shfl_xor_sync(0xffffffffu, v.s, o);
      v = ms_merge(v, other);
    }
    st = v;
  }
  return st;  // valid in warp 0
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

template <bool BF16>
__global__ void __launch_bounds__(kCeThreads)
ce_forward_kernel(const typename Elem<BF16>::T* __restrict__ logits,
                  const long long* __restrict__ labels, long long n_rows, int vocab,
                  long long row_stride, long long ignore_index, int split,
                  float* __restrict__ row_lse, float* __restrict__ loss_out,
                  float* __restrict__ nvalid_out, void* __restrict__ ws) {
  using E = Elem<BF16>;
  using T = typename E::T;
  constexpr int PP = E::kPerPack;
  constexpr int U = 4;
  __shared__ MS s_ms[kCeThreads / 32];
  __shared__ float s_f[kCeThreads / 32];
  __shared__ bool s_last;

  const CeScratch sc = carve(ws, n_rows);
  const long long row = blockIdx.x / split;
  const int part = (int)(blockIdx.x - row * split);
  const T* zrow = logits + row * row_stride;

  // Column range of this CTA, in whole packs when the row is 16-byte aligned.
  const bool vec_ok = ((((uintptr_t)zrow) & 15) == 0);
  const int npack = vec_ok ? vocab / PP : 0;
  const int per = (npack + split - 1) / split;
  const int p0 = min(part * per, npack), p1 = min(p0 + per, npack);

  MS st{-INFINITY, 0.f};
  const uint4* zp = reinterpret_cast<const uint4*>(zrow);
  int i = p0 + threadIdx.x;
  for (; i + (U - 1) * kCeThreads < p1; i += U * kCeThreads) {
    uint4 r[U];
#pragma unroll
    for (int u = 0; u < U; ++u) r[u] = ld_nc16(zp + i + u * kCeThreads);
    float x[U * PP];
#pragma unroll
    for (int u = 0; u < U; ++u) E::unpack(r[u], x + u * PP);
    ms_update<U * PP>(st, x);
  }
  for (; i < p1; i += kCeThreads) {
    float x[PP];
    E::unpack(ld_nc16(zp + i), x);
    ms_update<PP>(st, x);
  }
  // Scalar tail (vocab % PP, or the whole row if it is not 16-byte aligned): last part only.
  if (part == split - 1) {
    for (int j = npack * PP + threadIdx.x; j < vocab; j += kCeThreads) {
      float x[1] = {E::load1(zrow + j)};
      ms_update<1>(st, x);
    }
  }
  st = block_merge(st, s_ms);

  if (threadIdx.x == 0) {
    if (split == 1) {
      // Whole row seen by this CTA: finish the row here so the last CTA only has to sum.
      const float lse = st.m + logf(st.s);
      row_lse[row] = lse;
      const long long y = labels[row];
      float l = 0.f;
      if (y != ignore_index)
        l = (y < 0 || y >= vocab) ? __int_as_float(0x7fc00000)  // out-of-range label poisons the loss
                                  : lse - E::load1(zrow + y);
      sc.row_loss[row] = l;
    } else {
      sc.partial[row * split + part] = st;
    }
    __threadfence();
    const unsigned int arrived = atomicAdd(sc.counter, 1u);
    s_last = (arrived == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();

  // Last CTA to arrive: (split rows only) fold the partials of each row, then sum the valid
  // rows' losses in a fixed order (thread-strided, then a tree), so the result does not
  // depend on which CTA happened to be last.
  float acc = 0.f, cnt = 0.f;
  for (long long r = threadIdx.x; r < n_rows; r += kCeThreads) {
    const long long y = labels[r];
    const bool valid = (y != ignore_index);
    float l;
    if (split == 1) {
      l = __ldcg(sc.row_loss + r);
    } else {
      MS t{-INFINITY, 0.f};
      for (int k = 0; k < split; ++k) {
        const float2 q = __ldcg(reinterpret_cast<const float2*>(sc.partial + r * split + k));
        t = ms_merge(t, MS{q.x, q.y});
      }
      const float lse = t.m + logf(t.s);
      row_lse[r] = lse;
      l = 0.f;
      if (valid)
        l = (y < 0 || y >= vocab) ? __int_as_float(0x7fc00000)
                                  : lse - E::load1(logits + r * row_stride + y);
    }
    if (valid) {
      acc += l;
      cnt += 1.f;
    }
  }
  acc = block_sum(acc, s_f);
  cnt = block_sum(cnt, s_f);
  if (threadIdx.x == 0) {
    loss_out[0] = acc / cnt;  // 0/0 = NaN when every row is ignored, as torch
    nvalid_out[0] = cnt;
    *sc.counter = 0u;  // ready for the next call on this stream
  }
}

template <bool BF16>
__global__ void __launch_bounds__(kCeThreads)
ce_backward_kernel(const typename Elem<BF16>::T* __restrict__ logits,
                   const long long* __restrict__ labels, const float* __restrict__ row_lse,
                   const float* __restrict__ grad_out, const float* __restrict__ nvalid,
                   int vocab, long long row_stride, long long ignore_index, int tiles_per_row,
                   typename Elem<BF16>::T* __restrict__ dlogits) {
  using E = Elem<BF16>;
  using T = typename E::T;
  constexpr int PP = E::kPerPack;
  constexpr int U = 4;
  const long long row = blockIdx.x / tiles_per_row;
  const int tile = (int)(blockIdx.x - row * tiles_per_row);
  const T* zrow = logits + row * row_stride;
  T* drow = dlogits + row * row_stride;
  const long long y = labels[row];
  const bool valid = (y != ignore_index);
  const float coef = valid ? __ldg(grad_out) / __ldg(nvalid) : 0.f;
  const float lse = __ldg(row_lse + row);

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
        const int col0 = i * PP;
#pragma unroll
        for (int k = 0; k < PP; ++k) {
          const float p = fast_exp2((x[k] - lse) * kLog2e);
          x[k] = (p - ((long long)(col0 + k) == y ? 1.f : 0.f)) * coef;
        }
        dp[i] = E::pack(x);
      }
    }
  }
  if (tile == tiles_per_row - 1) {
    for (int j = npack * PP + threadIdx.x; j < vocab; j += kCeThreads) {
      float o = 0.f;
      if (valid) {
        const float p = fast_exp2((E::load1(zrow + j) - lse) * kLog2e);
        o = (p - ((long long)j == y ? 1.f : 0.f)) * coef;
      }
      E::store1(drow + j, o);
    }
  }
}

static int pick_split(long long n_rows) {
  // Split a row over several CTAs only while one CTA per row cannot fill the chip.
  const long long target = (long long)kNumSMs * 4;
  int s = 1;
  while (s < kCeMaxSplit && n_rows * s < target) s <<= 1;
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
  B2_REQUIRE(workspace_bytes >= b2_ce_workspace_bytes(n_rows, vocab),
             "ce_forward: workspace %zu B < required %zu B", workspace_bytes,
             b2_ce_workspace_bytes(n_rows, vocab));
  B2_REQUIRE((((uintptr_t)workspace) & 15) == 0, "ce_forward: workspace must be 16-byte aligned");
  const int split = pick_split(n_rows);
  const long long grid = n_rows * split;
  B2_REQUIRE(grid < (1ll << 31), "ce_forward: too many rows");
  if (dtype == B2_BF16) {
    ce_forward_kernel<true><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, workspace);
  } else {
    ce_forward_kernel<false><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
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
  const int per = 4 * kCeThreads;
  long long tiles = (npack + per - 1) / per;
  if (tiles < 1) tiles = 1;
  const long long grid = n_rows * tiles;
  B2_REQUIRE(grid < (1ll << 31), "ce_backward: too many tiles");
  if (dtype == B2_BF16) {
    ce_backward_kernel<true><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, row_lse, grad_out, nvalid,
        (int)vocab, row_stride, ignore_index, (int)tiles, (__nv_bfloat16*)dlogits);
  } else {
    ce_backward_kernel<false><<<(unsigned)grid, kCeThreads, 0, (cudaStream_t)stream>>>(
        (const float*)logits, (const long long*)labels, row_lse, grad_out, nvalid, (int)vocab,
        row_stride, ignore_index, (int)tiles, (float*)dlogits);
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
  const long long grid = n_rows * split;
  if (dtype == B2_BF16) {
    ce_forward_kernel<true><<<(unsigned)grid, kCeThreads, 0, st>>>(
        (const __nv_bfloat16*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, ws);
  } else {
    ce_forward_kernel<false><<<(unsigned)grid, kCeThreads, 0, st>>>(
        (const float*)logits, (const long long*)labels, n_rows, (int)vocab, row_stride,
        ignore_index, split, row_lse, loss_out, nvalid_out, ws);
  }
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}
