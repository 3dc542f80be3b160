// Global gradient norm + clip coefficient over many tensors in ONE launch.
//
// Replaces torch.nn.utils.clip_grad_norm_ as HF Trainer calls it between gradient sync and
// optimizer.step() (max_grad_norm=1.0 default: SP/transformers/training_args.py:857-859,
// SP/transformers/trainer.py:1753-1757; reached from train-torchrun.py:140). torch runs
// _foreach_norm + stack + norm + clamp + _foreach_mul_: two more full passes over the
// gradients. Here one read-only pass produces the norm and the clamped coefficient as DEVICE
// scalars, and the multiply is folded into the AdamW kernel (its grad_scale argument), so the
// gradients are read 2x instead of 3x and never rewritten.
//
// Deterministic: the chunk -> CTA assignment is static, every CTA reduces in a fixed order and
// the last CTA to arrive adds the per-CTA partials in index order.
// Roofline: HBM, 4 B per gradient element (read once).
#include "common.cuh"

namespace b2 {

constexpr int kGnThreads = 256;
constexpr int kGnCtasPerSM = 4;

__global__ void __launch_bounds__(kGnThreads, kGnCtasPerSM)
grad_sumsq_kernel(const b2_adamw_tensor* __restrict__ table, int n_tensors, long long total_chunks,
                  float max_norm, float* __restrict__ partials, unsigned int* __restrict__ counter,
                  float* __restrict__ norm_out, float* __restrict__ coef_out) {
  __shared__ int s_tensor;
  __shared__ float s_red[kGnThreads / 32];
  __shared__ bool s_last;
  float acc = 0.f;
  for (long long c = blockIdx.x; c < total_chunks; c += gridDim.x) {
    if (threadIdx.x == 0) {
      int lo = 0, hi = n_tensors - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((long long)__ldg(&table[mid].chunk_begin) <= c) lo = mid; else hi = mid - 1;
      }
      s_tensor = lo;
    }
    __syncthreads();
    const b2_adamw_tensor* e = table + s_tensor;
    __syncthreads();
    const long long begin = (c - e->chunk_begin) * (long long)B2_ADAMW_CHUNK;
    const long long left = e->numel - begin;
    const int count = left < B2_ADAMW_CHUNK ? (int)left : B2_ADAMW_CHUNK;
    const float* g = (const float*)e->g + begin;
    if ((((uintptr_t)g) & 15) == 0) {
      const int nvec = count >> 2;
      int i = threadIdx.x;
      for (; i + 3 * kGnThreads < nvec; i += 4 * kGnThreads) {
        uint4 r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) r[u] = ld_nc16(g + 4 * (i + u * kGnThreads));
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float x = __uint_as_float(r[u].x), y = __uint_as_float(r[u].y);
          const float z = __uint_as_float(r[u].z), w = __uint_as_float(r[u].w);
          acc += x * x + y * y + z * z + w * w;
        }
      }
      for (; i < nvec; i += kGnThreads) {
        const uint4 r = ld_nc16(g + 4 * i);
        const float x = __uint_as_float(r.x), y = __uint_as_float(r.y);
        const float z = __uint_as_float(r.z), w = __uint_as_float(r.w);
        acc += x * x + y * y + z * z + w * w;
      }
      const int tail = nvec << 2;
      if (threadIdx.x < count - tail) {
        const float x = g[tail + threadIdx.x];
        acc += x * x;
      }
    } else {
      for (int j = threadIdx.x; j < count; j += kGnThreads) acc += g[j] * g[j];
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < kGnThreads / 32 ? s_red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) {
      partials[blockIdx.x] = v;
      __threadfence();
      s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // fixed order: thread t adds partials t, t+256, ...; then a tree over the 256 threads (double
  // for the final few hundred terms costs nothing and keeps the norm within 1 ulp of exact)
  double total = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += kGnThreads) total += (double)__ldcg(partials + i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
  __shared__ double s_d[kGnThreads / 32];
  if ((threadIdx.x & 31) == 0) s_d[threadIdx.x >> 5] = total;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < kGnThreads / 32; ++i) t += s_d[i];
    const float norm = sqrtf((float)t);
    norm_out[0] = norm;
    // torch: clip_coef = max_norm / (total_norm + 1e-6), clamped to 1.0
    const float coef = max_norm / (norm + 1e-6f);
    coef_out[0] = coef > 1.0f ? 1.0f : coef;
    *counter = 0u;
  }
}

}  // namespace b2

extern "C" size_t b2_grad_norm_workspace_bytes(void) {
  return 16 + (size_t)b2::kNumSMs * b2::kGnCtasPerSM * sizeof(float);
}

extern "C" int b2_grad_norm_clip(const b2_adamw_tensor* tensors_dev, int32_t n_tensors,
                                 int64_t total_chunks, float max_norm, void* workspace,
                                 size_t workspace_bytes, float* norm_out, float* coef_out,
                                 b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(tensors_dev && workspace && norm_out && coef_out, "grad_norm_clip: null pointer");
  B2_REQUIRE(n_tensors > 0 && total_chunks > 0, "grad_norm_clip: empty table");
  B2_REQUIRE(max_norm > 0.f, "grad_norm_clip: max_norm must be positive");
  B2_REQUIRE(workspace_bytes >= b2_grad_norm_workspace_bytes() && (((uintptr_t)workspace) & 15) == 0,
             "grad_norm_clip: workspace too small or not 16-byte aligned");
  const long long full = (long long)kNumSMs * kGnCtasPerSM;
  const int grid = (int)(total_chunks < full ? total_chunks : full);
  grad_sumsq_kernel<<<grid, kGnThreads, 0, (cudaStream_t)stream>>>(
      tensors_dev, n_tensors, (long long)total_chunks, max_norm, (float*)((char*)workspace + 16),
      (unsigned int*)workspace, norm_out, coef_out);
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}
