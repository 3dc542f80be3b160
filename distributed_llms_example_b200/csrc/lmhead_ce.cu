// Fused lm_head + label-masked cross-entropy: the [N, V] logits never
// reach HBM in the forward pass.
//
// Replaces, for T5-family models, the pair
//     lm_logits = self.lm_head(sequence_output)                      (cuBLAS GEMM, writes N*V bf16)
//     loss = CrossEntropyLoss(ignore_index=-100)(lm_logits.view(-1, V), labels.view(-1))
// at SP/transformers/models/t5/modeling_t5.py:1110-1117, reached from train-accelerator.py:220-221,
// train-task.py:289-290 and train-torchrun.py:140. This is the one dense contraction on the hot
// path, so it runs on the 5th-generation tensor cores:
//
//   warp 0      TMA producer: cp.async.bulk.tensor (128B-swizzled boxes of hidden [128 x 64] and
//               weight [256 x 64] bf16) into a 4-stage shared-memory ring, mbarrier complete_tx
//   warp 1      one thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=256, K=16, bf16 x
//               bf16 -> fp32) on shared-memory descriptors; the accumulator lives in TMEM, double
//               buffered (2 x 256 columns = the whole 512-column TMEM), tcgen05.commit signals
//               "stage free" to the producer and "tile done" to the epilogue
//   warps 2-5   epilogue: tcgen05.ld 32 lanes x 32 columns per warp, ONE THREAD PER ROW, so the
//               online (max, sum-exp) and the label-logit pick are thread-local - no shuffles;
//               the next tile's MMAs run while this tile's 32768 exponentials are taken
//
// Work unit = (block of 128 rows, contiguous run of 256-wide vocabulary tiles); the run length is
// chosen on the host so that the grid fills 148 SMs in as few waves as possible. Each unit writes
// one (max, sum) partial per row; lmce_finish_kernel merges the partials per row in split order
// and reduces the loss with a fixed-order last-block sum (deterministic).
//
// Two more epilogues share the main loop:
//   forward + store   the bf16 logits are ALSO written, once, on their way out of TMEM (registers ->
//                     128B-swizzled shared memory -> cp.async.bulk.tensor store), so the forward
//                     never re-reads them and the backward can be the HBM-bound b2_ce_backward.
//                     This is the fastest pairing on B200: re-reading 2 bytes per logit costs less
//                     than recomputing 2*d_model flop per logit (measured, profiles/README.md).
//   dlogits           the logits tile is recomputed and the epilogue writes
//                       dlogits = (exp(z - lse) - [j == y]) * g / n_valid      (0 on ignored rows)
//                     in bf16: no [N, V] tensor is kept between forward and backward (memory-lean).
// dX = dlogits @ W and dW = dlogits^T @ X are plain library GEMMs in the Python adapter.
//
// Numerics: fp32 accumulation over K; with round_bf16 != 0 each logit is rounded to bf16 before
// the softmax, which is exactly what the unfused path computes from its bf16 logits tensor.
// Roofline: tensor pipe, 2*N*V*D flop per pass.
#include <cuda.h>

#include "common.cuh"

namespace b2 {
namespace lmce {

constexpr int kBM = 128;         // rows of hidden per tile (UMMA M)
constexpr int kBN = 256;         // vocabulary columns per tile (UMMA N)
constexpr int kBK = 64;          // K per stage: 64 bf16 = 128 bytes = one swizzle row
constexpr int kUmmaK = 16;       // K per tcgen05.mma for 16-bit inputs
constexpr int kStages = 4;
constexpr int kABytes = kBM * kBK * 2;   // 16 KiB
constexpr int kBBytes = kBN * kBK * 2;   // 32 KiB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kThreads = 192;    // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue
constexpr int kTmemCols = 512;
constexpr int kMaxSplits = 64;
constexpr int kGroup = 64;                 // columns per epilogue step: 64 bf16 = one 128-byte swizzle row
constexpr int kOutBytes = 32 * kGroup * 2; // per epilogue warp: 32 rows x 128 bytes staged for the TMA store
// shared memory: [stages x (A | B)] [4 x store staging] then barriers, 1024-byte aligned for the 128B swizzle
constexpr int kSmemBytes = kStages * kStageBytes + 4 * kOutBytes + 1024 /*alignment slack*/ + 256 /*barriers*/;

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// shared -> global tile store; rows / columns outside the tensor are clipped by the hardware
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"((uint64_t)map),
               "r"(smem_u32(src)), "r"(c0), "r"(c1)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)map) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, both operands K-major, issued by ONE thread for the CTA
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all MMAs issued so far by this thread arrive on the mbarrier when they complete
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// 32 TMEM lanes (this warp's quarter) x 32 consecutive fp32 columns -> 32 registers per thread
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for a K-major tile whose rows are 128 bytes (64 bf16), laid out
// by TMA with CU_TENSOR_MAP_SWIZZLE_128B: 8-row groups are 1024 bytes apart (stride byte offset),
// the leading byte offset is unused for swizzled K-major layouts, descriptor version 1 (sm_100),
// layout type 2 = SWIZZLE_128B. (cute/arch/mma_sm100_desc.hpp UMMA::SmemDescriptor.)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;            // leading byte offset (16 B units), ignored for SW128 K-major
  d |= (uint64_t)(1024 >> 4) << 32;  // stride byte offset: 8 rows x 128 B
  d |= (uint64_t)1 << 46;            // descriptor version
  d |= (uint64_t)2 << 61;            // SWIZZLE_128B
  return d;
}
// Instruction descriptor (UMMA::InstrDescriptor): D fp32, A/B bf16, both K-major, N=256, M=128.
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);

struct Barriers {
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

struct Params {
  const long long* labels;
  long long n_rows, vocab, d_model, ignore_index;
  int splits;            // vocabulary tile runs per row block
  int tiles_per_split;   // ceil(vocab tiles / splits)
  int n_vtiles;
  int round_bf16;
  // forward
  float2* partial;       // [n_rows, splits] (max in log2 units, sum)
  float* zlabel;         // [n_rows] label logit (written by the split that holds the label's column)
  // backward
  const float* row_lse;  // [n_rows], natural log
  const float* grad_out; // device scalar
  const float* nvalid;   // device scalar
};

__device__ __forceinline__ float round_to_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// One CTA = one (row block, split). MODE 0: forward (online LSE); MODE 1: backward (recompute,
// write dlogits); MODE 2: forward that also writes the bf16 logits.
template <int MODE>
__global__ void __launch_bounds__(kThreads, 1)
lmce_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
            const __grid_constant__ CUtensorMap map_out, const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* out_stage = smem + kStages * kStageBytes;
  Barriers* bars = (Barriers*)(out_stage + 4 * kOutBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rb = blockIdx.x / p.splits, split = blockIdx.x % p.splits;
  const int t_begin = split * p.tiles_per_split;
  const int t_end = min(t_begin + p.tiles_per_split, p.n_vtiles);
  const int n_tiles = max(t_end - t_begin, 0);
  const int n_kb = (int)(p.d_model / kBK);
  constexpr bool kStore = (MODE != 0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_w);
    if (kStore) tma_prefetch_desc(&map_out);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&bars->full[i], 1);
      mbar_init(&bars->empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->tmem_full[i], 1);
      mbar_init(&bars->tmem_empty[i], 4);  // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&bars->tmem_base, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = 0; t < n_tiles; ++t) {
        const int col0 = (t_begin + t) * kBN;
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(&bars->empty[stage], phase ^ 1);
          uint8_t* a = smem + stage * kStageBytes;
          mbar_expect_tx(&bars->full[stage], kStageBytes);
          tma_load_2d(&map_x, &bars->full[stage], a, kb * kBK, rb * kBM);
          tma_load_2d(&map_w, &bars->full[stage], a + kABytes, kb * kBK, col0);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = 0; t < n_tiles; ++t) {
        const int acc = t & 1;
        const uint32_t acc_phase = (uint32_t)(t >> 1) & 1;
        mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(acc * kBN);
        for (int kb = 0; kb < n_kb; ++kb) {
          mbar_wait(&bars->full[stage], phase);             // TMA bytes have landed
          tc_fence_after();
          const uint32_t a = smem_u32(smem + stage * kStageBytes);
          const uint64_t da = make_smem_desc(a), db = make_smem_desc(a + kABytes);
#pragma unroll
          for (int k = 0; k < kBK / kUmmaK; ++k) {
            // advance both descriptors by 16 bf16 = 32 bytes along K inside the 128-byte swizzle row
            umma_bf16(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), kIdesc, (uint32_t)((kb | k) != 0));
          }
          umma_commit(&bars->empty[stage]);                 // stage reusable once these MMAs retire
          if (kb == n_kb - 1) umma_commit(&bars->tmem_full[acc]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===== epilogue: thread <-> row =====
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const long long row = (long long)rb * kBM + q * 32 + lane;
    const bool row_ok = row < p.n_rows;
    long long label = row_ok ? p.labels[row] : p.ignore_index;
    const bool valid = row_ok && label != p.ignore_index;
    float m = -INFINITY, s = 0.f, zl = 0.f;
    bool have_label = false;
    float nlse2 = 0.f, coef = 0.f;
    if (MODE == 1) {
      nlse2 = row_ok ? -p.row_lse[row] * kLog2e : 0.f;
      coef = valid ? p.grad_out[0] / p.nvalid[0] : 0.f;
    }
    uint8_t* my_stage = out_stage + q * kOutBytes;   // 32 rows x 128 B, 128B-swizzled like the TMA box
    for (int t = 0; t < n_tiles; ++t) {
      const int acc = t & 1;
      const uint32_t acc_phase = (uint32_t)(t >> 1) & 1;
      const long long col0 = (long long)(t_begin + t) * kBN;
      mbar_wait(&bars->tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * kBN);
      const long long lc = valid ? label - col0 : -1;   // label's column inside this tile, if any
#pragma unroll 1
      for (int gi = 0; gi < kBN / kGroup; ++gi) {
        const long long cbase = col0 + gi * kGroup;
        if (cbase >= p.vocab) break;                    // whole group beyond the vocabulary (last tile)
        uint32_t v[kGroup];
        tmem_ld32(taddr + (uint32_t)(gi * kGroup), v);
        tmem_ld32(taddr + (uint32_t)(gi * kGroup + 32), v + 32);
        tmem_ld_wait();
        float z[kGroup];
#pragma unroll
        for (int j = 0; j < kGroup; ++j) {
          z[j] = __uint_as_float(v[j]);
          if (p.round_bf16) z[j] = round_to_bf16(z[j]);
        }
        const int ncol = (int)min((long long)kGroup, p.vocab - cbase);  // < 64 only in the last group
        const bool label_here = lc >= gi * kGroup && lc < gi * kGroup + kGroup;
        const int jl = (int)(lc - gi * kGroup);
        uint32_t o[kGroup / 2];
        if (MODE == 2) {
          if (ncol < kGroup) {
#pragma unroll
            for (int j = 0; j < kGroup; ++j)
              if (j >= ncol) z[j] = 0.f;   // padding columns of the last group are stored as zeros
          }
#pragma unroll
          for (int j = 0; j < kGroup; j += 2) o[j >> 1] = pack_bf16(z[j], z[j + 1]);
        }
        if (MODE != 1) {
          if (label_here) {
#pragma unroll
            for (int j = 0; j < kGroup; ++j)
              if (j == jl) zl = z[j];
            have_label = true;
          }
          float cm = -INFINITY;
#pragma unroll
          for (int j = 0; j < kGroup; ++j) {
            z[j] = (j < ncol) ? z[j] * kLog2e : -INFINITY;
            cm = fmaxf(cm, z[j]);
          }
          const float mn = fmaxf(m, cm);
          float add0 = 0.f, add1 = 0.f;
#pragma unroll
          for (int j = 0; j < kGroup; j += 2) {
            add0 += fast_exp2(z[j] - mn);
            add1 += fast_exp2(z[j + 1] - mn);
          }
          s = s * fast_exp2(m - mn) + (add0 + add1);
          m = mn;
        } else {
          // dlogits = (softmax - onehot) * g / n_valid
#pragma unroll
          for (int j = 0; j < kGroup; ++j) z[j] = fast_exp2(__fmaf_rn(z[j], kLog2e, nlse2)) * coef;
          if (label_here) {
#pragma unroll
            for (int j = 0; j < kGroup; ++j)
              if (j == jl) z[j] -= coef;
          }
          if (ncol < kGroup) {
#pragma unroll
            for (int j = 0; j < kGroup; ++j)
              if (j >= ncol) z[j] = 0.f;
          }
#pragma unroll
          for (int j = 0; j < kGroup; j += 2) o[j >> 1] = pack_bf16(z[j], z[j + 1]);
        }
        if (kStore) {
          // registers -> swizzled shared memory -> one TMA store of [32 rows x 64 columns] per warp.
          // The previous store must have finished READING the staging tile before it is overwritten.
          if (lane == 0) tma_store_wait_read();
          __syncwarp();
          uint8_t* rowp = my_stage + lane * 128;
#pragma unroll
          for (int c16 = 0; c16 < 8; ++c16)
            *reinterpret_cast<uint4*>(rowp + ((c16 ^ (lane & 7)) << 4)) =
                make_uint4(o[4 * c16], o[4 * c16 + 1], o[4 * c16 + 2], o[4 * c16 + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) tma_store_2d(&map_out, my_stage, (int)cbase, (int)((long long)rb * kBM + q * 32));
        }
      }
      // this accumulator may be overwritten by the MMA of tile t+2
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
    }
    if (MODE != 1 && row_ok) {
      p.partial[row * p.splits + split] = make_float2(m, s);
      if (have_label) p.zlabel[row] = zl;
    }
    if (kStore && lane == 0) tma_store_wait_all();   // bulk stores complete before the CTA may exit
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// Merge the per-split (max, sum) partials of each row, in split order; then a deterministic
// last-block reduction of sum(lse - z_label) over valid rows and of the valid count.
constexpr int kFinThreads = 256;
__global__ void __launch_bounds__(kFinThreads)
lmce_finish_kernel(const float2* __restrict__ partial, const float* __restrict__ zlabel,
                   const long long* __restrict__ labels, long long n_rows, long long vocab, int splits,
                   long long ignore_index, float* __restrict__ row_lse, float* __restrict__ block_sums,
                   unsigned int* __restrict__ counter, float* __restrict__ loss_out, float* __restrict__ nvalid_out) {
  __shared__ float s_loss[kFinThreads / 32], s_cnt[kFinThreads / 32];
  __shared__ bool s_last;
  const long long row = (long long)blockIdx.x * kFinThreads + threadIdx.x;
  float loss = 0.f, cnt = 0.f;
  if (row < n_rows) {
    float m = -INFINITY;
    for (int i = 0; i < splits; ++i) m = fmaxf(m, partial[row * splits + i].x);
    float s = 0.f;
    for (int i = 0; i < splits; ++i) {
      const float2 pr = partial[row * splits + i];
      s += pr.y * fast_exp2(pr.x - m);
    }
    const float lse = (m + __log2f(s)) * kLn2;
    row_lse[row] = lse;
    const long long y = labels[row];
    if (y != ignore_index) {
      // like b2_ce_forward: a label outside [0, vocab) poisons the loss with NaN
      loss = (y >= 0 && y < vocab) ? lse - zlabel[row] : __int_as_float(0x7fc00000);
      cnt = 1.f;
    }
  }
  loss = warp_sum(loss);
  cnt = warp_sum(cnt);
  if ((threadIdx.x & 31) == 0) {
    s_loss[threadIdx.x >> 5] = loss;
    s_cnt[threadIdx.x >> 5] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, c = 0.f;
    for (int i = 0; i < kFinThreads / 32; ++i) { l += s_loss[i]; c += s_cnt[i]; }
    block_sums[2 * blockIdx.x] = l;
    block_sums[2 * blockIdx.x + 1] = c;
    __threadfence();
    s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last || threadIdx.x != 0) return;
  __threadfence();
  double l = 0.0, c = 0.0;
  for (unsigned i = 0; i < gridDim.x; ++i) {
    l += (double)__ldcg(block_sums + 2 * i);
    c += (double)__ldcg(block_sums + 2 * i + 1);
  }
  nvalid_out[0] = (float)c;
  loss_out[0] = (float)(l / c);  // 0/0 = NaN when no row is valid, like torch
  *counter = 0u;
}

// ---- host side ---------------------------------------------------------------------------------
typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_fn(EncodeTiled* out) {
  static EncodeTiled fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    B2_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q));
    if (!f || q != cudaDriverEntryPointSuccess)
      return fail(B2_ERR_CUDA, "lmce: cuTensorMapEncodeTiled is not available from this driver");
    fn = (EncodeTiled)f;
  }
  *out = fn;
  return B2_OK;
}

// [rows, d_model] bf16 row-major -> boxes of [box_rows, 64] with the 128-byte swizzle; rows beyond
// the tensor read as zeros (they only ever feed masked logits)
static int make_map(CUtensorMap* map, const void* base, long long rows, long long d_model, int box_rows) {
  EncodeTiled enc;
  int rc = encode_fn(&enc);
  if (rc != B2_OK) return rc;
  const cuuint64_t dims[2] = {(cuuint64_t)d_model, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)d_model * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(B2_ERR_CUDA, "lmce: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return B2_OK;
}

// [rows, vocab] bf16 output with row pitch ld: boxes of [32 rows, 64 columns], 128-byte swizzle
static int make_out_map(CUtensorMap* map, void* base, long long rows, long long vocab, long long ld) {
  EncodeTiled enc;
  int rc = encode_fn(&enc);
  if (rc != B2_OK) return rc;
  // stores move whole 16-byte units: the columns [vocab, vocab rounded up to 8) receive zeros
  const cuuint64_t dims[2] = {(cuuint64_t)((vocab + 7) / 8 * 8), (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kGroup, 32};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(B2_ERR_CUDA, "lmce: cuTensorMapEncodeTiled (output) failed with CUresult %d", (int)r);
  return B2_OK;
}

// splits per row block: fewest (waves x tiles per CTA), ties to the smaller partial array
static int pick_splits(long long row_blocks, int n_vtiles) {
  long long best_cost = -1;
  int best = 1;
  for (int s = 1; s <= kMaxSplits && s <= n_vtiles; ++s) {
    const long long ctas = row_blocks * s;
    const long long waves = (ctas + kNumSMs - 1) / kNumSMs;
    const long long tiles = (n_vtiles + s - 1) / s;
    const long long cost = waves * tiles;
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = s;
    }
  }
  return best;
}

static int check_common(const void* hidden, const void* weight, const long long* labels, long long n_rows,
                        long long vocab, long long d_model) {
  B2_REQUIRE(hidden && weight && labels, "lmce: null pointer");
  B2_REQUIRE(n_rows > 0 && vocab > 0, "lmce: empty problem (%lld rows, vocab %lld)", n_rows, vocab);
  B2_REQUIRE(d_model >= kBK && d_model % kBK == 0, "lmce: d_model %lld must be a positive multiple of %d", d_model, kBK);
  B2_REQUIRE(((((uintptr_t)hidden) | ((uintptr_t)weight)) & 15) == 0, "lmce: hidden / weight must be 16-byte aligned");
  return B2_OK;
}

static int set_smem_attr(const void* fn) {
  B2_CUDA_CHECK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  return B2_OK;
}

}  // namespace lmce
}  // namespace b2

extern "C" size_t b2_lmce_workspace_bytes(int64_t n_rows, int64_t vocab) {
  using namespace b2::lmce;
  if (n_rows <= 0 || vocab <= 0) return 0;
  const long long blocks = (n_rows + kFinThreads - 1) / kFinThreads;
  // [counter, pad] | partial [n_rows, kMaxSplits] float2 | zlabel [n_rows] | block sums [blocks, 2]
  return 16 + (size_t)n_rows * kMaxSplits * 8 + (size_t)n_rows * 4 + (size_t)blocks * 8 + 64;
}

static int fill_params(b2::lmce::Params& p, const int64_t* labels, int64_t n_rows, int64_t vocab, int64_t d_model,
                       int64_t ignore_index, int round_bf16) {
  using namespace b2::lmce;
  p = Params{};
  p.labels = (const long long*)labels;
  p.n_rows = n_rows; p.vocab = vocab; p.d_model = d_model; p.ignore_index = ignore_index;
  p.n_vtiles = (int)((vocab + kBN - 1) / kBN);
  const long long row_blocks = (n_rows + kBM - 1) / kBM;
  p.splits = pick_splits(row_blocks, p.n_vtiles);
  p.tiles_per_split = (p.n_vtiles + p.splits - 1) / p.splits;
  p.splits = (p.n_vtiles + p.tiles_per_split - 1) / p.tiles_per_split;  // no empty trailing split
  p.round_bf16 = round_bf16;
  return (int)(row_blocks * p.splits);
}

extern "C" int b2_lmce_forward(const void* hidden, const void* weight, const int64_t* labels, int64_t n_rows,
                               int64_t vocab, int64_t d_model, int64_t ignore_index, int round_bf16,
                               float* row_lse, float* loss_out, float* nvalid_out, void* logits_out,
                               int64_t logits_stride, void* workspace, size_t workspace_bytes,
                               b2_stream_t stream) {
  using namespace b2;
  using namespace b2::lmce;
  int rc = check_common(hidden, weight, (const long long*)labels, n_rows, vocab, d_model);
  if (rc != B2_OK) return rc;
  B2_REQUIRE(row_lse && loss_out && nvalid_out && workspace, "lmce_forward: null pointer");
  B2_REQUIRE(workspace_bytes >= b2_lmce_workspace_bytes(n_rows, vocab) && (((uintptr_t)workspace) & 15) == 0,
             "lmce_forward: workspace too small or not 16-byte aligned");
  B2_REQUIRE(!logits_out || (logits_stride >= vocab && logits_stride % 8 == 0 && (((uintptr_t)logits_out) & 15) == 0),
             "lmce_forward: logits_out needs a 16-byte aligned base and a row pitch >= vocab that is a multiple of 8");
  CUtensorMap map_x, map_w, map_out;
  if ((rc = make_map(&map_x, hidden, n_rows, d_model, kBM)) != B2_OK) return rc;
  if ((rc = make_map(&map_w, weight, vocab, d_model, kBN)) != B2_OK) return rc;
  if (logits_out) {
    if ((rc = make_out_map(&map_out, logits_out, n_rows, vocab, logits_stride)) != B2_OK) return rc;
  } else {
    map_out = map_x;  // unused by the kernel variant launched below
  }
  Params p;
  const int grid = fill_params(p, labels, n_rows, vocab, d_model, ignore_index, round_bf16);
  char* ws = (char*)workspace;
  unsigned int* counter = (unsigned int*)ws;
  p.partial = (float2*)(ws + 16);
  p.zlabel = (float*)(ws + 16 + (size_t)n_rows * kMaxSplits * 8);
  float* block_sums = p.zlabel + n_rows;
  static bool attr_done = false;
  if (!attr_done) {
    if ((rc = set_smem_attr((const void*)lmce_kernel<0>)) != B2_OK) return rc;
    if ((rc = set_smem_attr((const void*)lmce_kernel<2>)) != B2_OK) return rc;
    attr_done = true;
  }
  if (logits_out)
    lmce_kernel<2><<<(unsigned)grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(map_x, map_w, map_out, p);
  else
    lmce_kernel<0><<<(unsigned)grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(map_x, map_w, map_out, p);
  B2_CUDA_CHECK(cudaGetLastError());
  const unsigned fin_blocks = (unsigned)((n_rows + kFinThreads - 1) / kFinThreads);
  lmce_finish_kernel<<<fin_blocks, kFinThreads, 0, (cudaStream_t)stream>>>(
      p.partial, p.zlabel, p.labels, n_rows, vocab, p.splits, ignore_index, row_lse, block_sums, counter, loss_out,
      nvalid_out);
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_lmce_backward_dlogits(const void* hidden, const void* weight, const int64_t* labels,
                                        const float* row_lse, const float* grad_out, const float* nvalid,
                                        int64_t n_rows, int64_t vocab, int64_t d_model, int64_t ignore_index,
                                        int round_bf16, void* dlogits, int64_t row_stride, b2_stream_t stream) {
  using namespace b2;
  using namespace b2::lmce;
  int rc = check_common(hidden, weight, (const long long*)labels, n_rows, vocab, d_model);
  if (rc != B2_OK) return rc;
  B2_REQUIRE(row_lse && grad_out && nvalid && dlogits, "lmce_backward: null pointer");
  B2_REQUIRE(row_stride >= vocab && row_stride % 8 == 0 && (((uintptr_t)dlogits) & 15) == 0,
             "lmce_backward: dlogits needs a 16-byte aligned base and a row pitch >= vocab that is a multiple of 8");
  CUtensorMap map_x, map_w, map_out;
  if ((rc = make_map(&map_x, hidden, n_rows, d_model, kBM)) != B2_OK) return rc;
  if ((rc = make_map(&map_w, weight, vocab, d_model, kBN)) != B2_OK) return rc;
  if ((rc = make_out_map(&map_out, dlogits, n_rows, vocab, row_stride)) != B2_OK) return rc;
  Params p;
  const int grid = fill_params(p, labels, n_rows, vocab, d_model, ignore_index, round_bf16);
  p.row_lse = row_lse; p.grad_out = grad_out; p.nvalid = nvalid;
  static bool attr_done = false;
  if (!attr_done) {
    if ((rc = set_smem_attr((const void*)lmce_kernel<1>)) != B2_OK) return rc;
    attr_done = true;
  }
  lmce_kernel<1><<<(unsigned)grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(map_x, map_w, map_out, p);
  B2_CUDA_CHECK(cudaGetLastError());
  return B2_OK;
}
