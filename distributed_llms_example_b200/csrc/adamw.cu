// AdamW step over many tensors in ONE launch (include/b2ddp.h, section (b)).
//
// Replaces torch.optim.AdamW.step as called at train-accelerator.py:224, train-task.py:295
// and by HF Trainer for train-torchrun.py:140. The arithmetic follows torch's single-tensor
// path operation by operation (SP/torch/optim/adam.py:436-536: mul_ decay, lerp_, mul_ +
// addcmul_, sqrt / bias_correction2_sqrt + eps, addcdiv_) so that results agree with the CPU
// oracle to the last bit in the common case.
//
// Roofline: pure HBM streaming, 28 B per element (read p,g,m,v; write p,m,v). Layout: the
// tensors stay where torch allocated them (131-557 separate allocations); a device table of
// {p,g,m,v,numel,hyper,chunk_begin} rows maps a flat work-item index (8192 elements) to its
// tensor by binary search, so one persistent grid of 148*k CTAs covers every tensor without
// per-tensor launches and without re-homing parameters.
#include <cmath>

#include "common.cuh"

namespace b2 {

struct DevHyper {
  float decay;          // 1 - lr*wd
  float w1;             // 1 - beta1 (lerp weight)
  float beta2;
  float omb2;           // 1 - beta2
  float neg_step_size;  // -(lr / bias_correction1)
  float bc2_sqrt;       // sqrt(bias_correction2)
  float eps;
  float pad;
};

struct HyperPack {
  DevHyper h[B2_ADAMW_MAX_HYPERS];
};

constexpr int kThreads = 256;
constexpr int kCtasPerSM = 3;

__device__ __forceinline__ void adamw_elem(float& p, float g, float& m, float& v,
                                           const DevHyper& h) {
  // Explicit _rn intrinsics keep nvcc from contracting into FMAs that torch's CPU
  // vectorised kernels do not use; the lerp IS an fma there (vec::fmadd).
  p = __fmul_rn(p, h.decay);
  m = __fmaf_rn(h.w1, __fsub_rn(g, m), m);
  v = __fadd_rn(__fmul_rn(v, h.beta2), __fmul_rn(__fmul_rn(h.omb2, g), g));
  float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), h.bc2_sqrt), h.eps);
  p = __fadd_rn(p, __fdiv_rn(__fmul_rn(h.neg_step_size, m), denom));
}

__device__ __forceinline__ void adamw_vec4(float4& p, float4 g, float4& m, float4& v,
                                           const DevHyper& h, float gs) {
  if (gs != 1.0f) {
    g.x *= gs; g.y *= gs; g.z *= gs; g.w *= gs;
  }
  adamw_elem(p.x, g.x, m.x, v.x, h);
  adamw_elem(p.y, g.y, m.y, v.y, h);
  adamw_elem(p.z, g.z, m.z, v.z, h);
  adamw_elem(p.w, g.w, m.w, v.w, h);
}

__device__ __forceinline__ float4 as_f4(uint4 u) {
  return make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z),
                     __uint_as_float(u.w));
}
__device__ __forceinline__ uint4 as_u4(float4 f) {
  return make_uint4(__float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(f.z),
                    __float_as_uint(f.w));
}

// One work item: elements [begin, begin+count) of one tensor.
__device__ __forceinline__ void adamw_chunk(float* __restrict__ p, const float* __restrict__ g,
                                            float* __restrict__ m, float* __restrict__ v,
                                            int count, const DevHyper& h, float gs) {
  const bool aligned = (((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0;
  if (aligned) {
    const int nvec = count >> 2;
    constexpr int U = 2;  // 2 x 4 streams x 16 B = 128 B in flight per thread
    int i = threadIdx.x;
    for (; i + (U - 1) * kThreads < nvec; i += U * kThreads) {
      uint4 rp[U], rg[U], rm[U], rv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int j = i + u * kThreads;
        rp[u] = ld_stream16(p + 4 * j);
        rg[u] = ld_nc16(g + 4 * j);
        rm[u] = ld_stream16(m + 4 * j);
        rv[u] = ld_stream16(v + 4 * j);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int j = i + u * kThreads;
        float4 fp = as_f4(rp[u]), fm = as_f4(rm[u]), fv = as_f4(rv[u]);
        adamw_vec4(fp, as_f4(rg[u]), fm, fv, h, gs);
        st_stream16(p + 4 * j, as_u4(fp));
        st_stream16(m + 4 * j, as_u4(fm));
        st_stream16(v + 4 * j, as_u4(fv));
      }
    }
    for (; i < nvec; i += kThreads) {
      float4 fp = as_f4(ld_stream16(p + 4 * i)), fm = as_f4(ld_stream16(m + 4 * i)),
             fv = as_f4(ld_stream16(v + 4 * i));
      adamw_vec4(fp, as_f4(ld_nc16(g + 4 * i)), fm, fv, h, gs);
      st_stream16(p + 4 * i, as_u4(fp));
      st_stream16(m + 4 * i, as_u4(fm));
      st_stream16(v + 4 * i, as_u4(fv));
    }
    const int tail = nvec << 2;
    if (threadIdx.x < count - tail) {
      const int j = tail + threadIdx.x;
      float fp = p[j], fm = m[j], fv = v[j];
      adamw_elem(fp, g[j] * gs, fm, fv, h);
      p[j] = fp; m[j] = fm; v[j] = fv;
    }
  } else {
    for (int j = threadIdx.x; j < count; j += kThreads) {
      float fp = p[j], fm = m[j], fv = v[j];
      adamw_elem(fp, g[j] * gs, fm, fv, h);
      p[j] = fp; m[j] = fm; v[j] = fv;
    }
  }
}

__global__ void __launch_bounds__(kThreads, kCtasPerSM)
adamw_multi_kernel(const b2_adamw_tensor* __restrict__ table, int n_tensors,
                   long long total_chunks, const __grid_constant__ HyperPack hp,
                   const float* __restrict__ grad_scale) {
  const float gs = grad_scale ? __ldg(grad_scale) : 1.0f;
  __shared__ int s_tensor;
  for (long long c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    // Binary search: last tensor whose chunk_begin <= c. Warp 0 lane 0 does it; the
    // table (<= 27 KB for flan-t5-large) is L1/L2 resident after the first pass.
    if (threadIdx.x == 0) {
      int lo = 0, hi = n_tensors - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((long long)__ldg(&table[mid].chunk_begin) <= c) lo = mid; else hi = mid - 1;
      }
      s_tensor = lo;
    }
    __syncthreads();
    const int t = s_tensor;
    __syncthreads();
    const b2_adamw_tensor* e = table + t;
    const long long numel = e->numel;
    const long long begin = (c - e->chunk_begin) * (long long)B2_ADAMW_CHUNK;
    const long long left = numel - begin;
    const int count = left < B2_ADAMW_CHUNK ? (int)left : B2_ADAMW_CHUNK;
    adamw_chunk((float*)e->p + begin, (const float*)e->g + begin, (float*)e->m + begin,
                (float*)e->v + begin, count, hp.h[e->hyper], gs);
  }
}

__global__ void __launch_bounds__(kThreads, kCtasPerSM)
adamw_flat_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                  float* __restrict__ v, long long numel, const DevHyper h,
                  const float* __restrict__ grad_scale) {
  const float gs = grad_scale ? __ldg(grad_scale) : 1.0f;
  const long long total_chunks = (numel + B2_ADAMW_CHUNK - 1) / B2_ADAMW_CHUNK;
  for (long long c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    const long long begin = c * (long long)B2_ADAMW_CHUNK;
    const long long left = numel - begin;
    const int count = left < B2_ADAMW_CHUNK ? (int)left : B2_ADAMW_CHUNK;
    adamw_chunk(p + begin, g + begin, m + begin, v + begin, count, h, gs);
  }
}

static int derive(const b2_adamw_hyper& in, DevHyper* out) {
  if (!(in.step >= 1)) return fail(B2_ERR_INVALID, "adamw: step must be >= 1 (got %lld)", (long long)in.step);
  if (!(in.lr >= 0.0) || !(in.eps >= 0.0) || !(in.beta1 >= 0.0 && in.beta1 < 1.0) ||
      !(in.beta2 >= 0.0 && in.beta2 < 1.0) || !(in.weight_decay >= 0.0))
    return fail(B2_ERR_INVALID, "adamw: invalid hyper-parameters (lr=%g betas=(%g,%g) eps=%g wd=%g)",
                in.lr, in.beta1, in.beta2, in.eps, in.weight_decay);
  // Same expressions, in double, as SP/torch/optim/adam.py:436,499-500,529-534.
  const double bc1 = 1.0 - std::pow(in.beta1, (double)in.step);
  const double bc2 = 1.0 - std::pow(in.beta2, (double)in.step);
  out->decay = (float)(1.0 - in.lr * in.weight_decay);
  out->w1 = (float)(1.0 - in.beta1);
  out->beta2 = (float)in.beta2;
  out->omb2 = (float)(1.0 - in.beta2);
  out->neg_step_size = (float)(-(in.lr / bc1));
  out->bc2_sqrt = (float)std::sqrt(bc2);
  out->eps = (float)in.eps;
  out->pad = 0.f;
  return B2_OK;
}

static int grid_for(long long total_chunks) {
  const long long full = (long long)kNumSMs * kCtasPerSM;
  return (int)(total_chunks < full ? total_chunks : full);
}

}  // namespace b2

extern "C" int64_t b2_adamw_plan(b2_adamw_tensor* tensors_host, int32_t n_tensors) {
  if (!tensors_host || n_tensors < 0) return b2::fail(B2_ERR_INVALID, "adamw_plan: null table");
  int64_t chunks = 0;
  for (int32_t i = 0; i < n_tensors; ++i) {
    b2_adamw_tensor& t = tensors_host[i];
    if (t.numel <= 0) return b2::fail(B2_ERR_INVALID, "adamw_plan: tensor %d has numel %lld", i, (long long)t.numel);
    if (t.hyper < 0 || t.hyper >= B2_ADAMW_MAX_HYPERS)
      return b2::fail(B2_ERR_INVALID, "adamw_plan: tensor %d hyper index %d out of range", i, t.hyper);
    if (chunks > INT32_MAX) return b2::fail(B2_ERR_TOO_LARGE, "adamw_plan: more than 2^31 work items");
    t.chunk_begin = (int32_t)chunks;
    chunks += (t.numel + B2_ADAMW_CHUNK - 1) / B2_ADAMW_CHUNK;
  }
  if (chunks > INT32_MAX) return b2::fail(B2_ERR_TOO_LARGE, "adamw_plan: more than 2^31 work items");
  return chunks;
}

extern "C" int b2_adamw_step(const b2_adamw_tensor* tensors_dev, int32_t n_tensors,
                             int64_t total_chunks, const b2_adamw_hyper* hypers_host,
                             int32_t n_hypers, const float* grad_scale, b2_stream_t stream) {
  using namespace b2;
  if (n_tensors == 0 || total_chunks == 0) return B2_OK;
  B2_REQUIRE(tensors_dev && hypers_host, "adamw_step: null pointer");
  B2_REQUIRE(n_tensors > 0 && total_chunks > 0, "adamw_step: negative sizes");
  B2_REQUIRE(n_hypers >= 1 && n_hypers <= B2_ADAMW_MAX_HYPERS,
             "adamw_step: n_hypers %d not in [1, %d]", n_hypers, B2_ADAMW_MAX_HYPERS);
  HyperPack hp;
  for (int i = 0; i < B2_ADAMW_MAX_HYPERS; ++i) {
    if (i < n_hypers) {
      int rc = derive(hypers_host[i], &hp.h[i]);
      if (rc != B2_OK) return rc;
    } else {
      hp.h[i] = hp.h[0];
    }
  }
  adamw_multi_kernel<<<grid_for(total_chunks), kThreads, 0, (cudaStream_t)stream>>>(
      tensors_dev, n_tensors, (long long)total_chunks, hp, grad_scale);
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_adamw_step_flat(void* p, const void* g, void* m, void* v, int64_t numel,
                                  const b2_adamw_hyper* hyper_host, const float* grad_scale,
                                  b2_stream_t stream) {
  using namespace b2;
  if (numel == 0) return B2_OK;
  B2_REQUIRE(p && g && m && v && hyper_host, "adamw_step_flat: null pointer");
  B2_REQUIRE(numel > 0, "adamw_step_flat: negative numel");
  DevHyper h;
  int rc = derive(*hyper_host, &h);
  if (rc != B2_OK) return rc;
  const long long chunks = (numel + B2_ADAMW_CHUNK - 1) / B2_ADAMW_CHUNK;
  adamw_flat_kernel<<<grid_for(chunks), kThreads, 0, (cudaStream_t)stream>>>(
      (float*)p, (const float*)g, (float*)m, (float*)v, (long long)numel, h, grad_scale);
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}
