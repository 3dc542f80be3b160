// Gradient all-reduce with in-switch reduction (NVLink SHARP / NVLS), SURVEY.md 8f row 3.
//
// Plain peer-to-peer all-reduce has to move 2(W-1)/W*S through every GPU's NVLink port in EACH
// direction. With a multicast object bound to one buffer per GPU, `multimem.ld_reduce` asks the
// NVSwitch for "the sum over all GPUs of the word at this offset" and `multimem.st` stores one
// word into every GPU's buffer: the owner of a slice pulls S/W reduced bytes and pushes S/W,
// and each port carries ~S per direction instead of 1.75*S at W=8.
//
//   one kernel per call (all W ranks):
//     copy-in   bucket -> this rank's multicast-bound buffer (unicast mapping), x scale
//     barrier   per-CTA flags in the communicator's IPC-mapped pad (allreduce.cu)
//     reduce    slice owner: v = multimem.ld_reduce.add.f32x4 [mc + i]; multimem.st [mc + i], v
//     barrier
//     copy-out  this rank's buffer -> bucket
//
// Every element is reduced exactly once (by its slice owner, inside the switch) and broadcast,
// so all ranks hold bit-identical results; the order in which the switch adds the W values is
// the hardware's, so against the rank-ordered oracle this path is equal to fp32 rounding
// (<= 1e-6 relative), not bit for bit - the same standard as NCCL's own ring/tree/NVLS orders.
// A single buffer suffices across consecutive calls: a rank's next copy-in only overwrites its
// own buffer after it passed the second barrier of this call, which every peer enters only after
// finishing its switch reads.
//
// Set-up (host, all through driver entry points resolved at run time, no libcuda link):
//   rank 0: cuMulticastCreate -> POSIX fd, passed to the other ranks by the host (SCM_RIGHTS);
//   all: cuMulticastAddDevice; barrier; cuMemCreate + cuMulticastBindMem; map unicast and
//   multicast virtual ranges.
#include <cuda.h>
#include <unistd.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <utility>
#include <vector>

#include "allreduce_shared.cuh"

namespace b2 {

// communicator whose multicast-bound region backs the torch MemPool (one per process), and the
// bump pointer inside it; used by b2_pool_alloc / b2_pool_free at the end of this file
static b2_comm* g_pool_comm = nullptr;
static size_t g_pool_next = 0;
static std::vector<std::pair<size_t, size_t>> g_pool_free;  // (offset, bytes), sorted by offset, coalesced
static std::mutex g_pool_mutex;

struct Drv {
  CUresult (*memGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*memCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*memRelease)(CUmemGenericAllocationHandle);
  CUresult (*memAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*memAddressFree)(CUdeviceptr, size_t);
  CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*memUnmap)(CUdeviceptr, size_t);
  CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*memExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*memImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*multicastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*multicastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*multicastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*multicastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*multicastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*deviceGet)(CUdevice*, int);
  bool ok = false;
};

static int load_one(const char* name, void** out) {
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, out, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !*out)
    return fail(B2_ERR_CUDA, "nvls: driver entry point %s is not available", name);
  return B2_OK;
}

static int drv(Drv** out) {
  static Drv d;
  if (!d.ok) {
#define B2_LOAD(field, name)                                  \
  {                                                           \
    int rc = load_one(name, (void**)&d.field);                \
    if (rc != B2_OK) return rc;                               \
  }
    B2_LOAD(memGetAllocationGranularity, "cuMemGetAllocationGranularity")
    B2_LOAD(memCreate, "cuMemCreate")
    B2_LOAD(memRelease, "cuMemRelease")
    B2_LOAD(memAddressReserve, "cuMemAddressReserve")
    B2_LOAD(memAddressFree, "cuMemAddressFree")
    B2_LOAD(memMap, "cuMemMap")
    B2_LOAD(memUnmap, "cuMemUnmap")
    B2_LOAD(memSetAccess, "cuMemSetAccess")
    B2_LOAD(memExportToShareableHandle, "cuMemExportToShareableHandle")
    B2_LOAD(memImportFromShareableHandle, "cuMemImportFromShareableHandle")
    B2_LOAD(multicastCreate, "cuMulticastCreate")
    B2_LOAD(multicastAddDevice, "cuMulticastAddDevice")
    B2_LOAD(multicastBindMem, "cuMulticastBindMem")
    B2_LOAD(multicastGetGranularity, "cuMulticastGetGranularity")
    B2_LOAD(multicastUnbind, "cuMulticastUnbind")
    B2_LOAD(deviceGet, "cuDeviceGet")
#undef B2_LOAD
    d.ok = true;
  }
  *out = &d;
  return B2_OK;
}

#define B2_DRV_CHECK(expr)                                                                    \
  do {                                                                                        \
    CUresult _r = (expr);                                                                     \
    if (_r != CUDA_SUCCESS)                                                                   \
      return ::b2::fail(B2_ERR_CUDA, "%s failed with CUresult %d (%s:%d)", #expr, (int)_r,    \
                        __FILE__, __LINE__);                                                  \
  } while (0)

// ---- device side ----------------------------------------------------------------------------
__device__ __forceinline__ uint4 multimem_ld_reduce_f32x4(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_f32x4(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ float multimem_ld_reduce_f32(const void* mc) {
  float r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(r) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_f32(void* mc, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}

struct NvlsArgs {
  RankView v;   // buf = bucket; flags[] = communicator pads; stage[] unused
  char* uc;     // this rank's multicast-bound buffer, unicast mapping
  char* mc;     // the multicast mapping of all ranks' buffers
};

template <int W>
__global__ void __launch_bounds__(kArThreads, 1)
allreduce_nvls_kernel(const __grid_constant__ NvlsArgs a, long long numel, float scale, uint32_t epoch,
                      unsigned long long timeout_ns) {
  const RankView& v = a.v;
  const int rank = v.rank;
  const int nb = gridDim.x, b = blockIdx.x, tid = threadIdx.x;
  const long long npack = numel / 4;
  const long long tail0 = npack * 4;
  float* __restrict__ buf = v.buf;
  const long long Q = (npack + W - 1) / W;
  const long long q = (Q + nb - 1) / nb;
  constexpr int U = 4;
  float sq = 0.f;

  if (comm_poisoned(v)) { poison_local(v, numel); return; }
  // copy-in: this CTA stages its sub-range of every slice (the peer CTA with the same index
  // is the one whose switch read will touch it after the per-CTA barrier)
#pragma unroll 1
  for (int s = 0; s < W; ++s) {
    const long long slo = min((long long)s * Q, npack), shi = min(slo + Q, npack);
    const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
    long long i = lo + tid;
    for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
      uint4 t[U];
#pragma unroll
      for (int u = 0; u < U; ++u) t[u] = Wire<false>::load_scaled(buf, i + u * kArThreads, scale);
#pragma unroll
      for (int u = 0; u < U; ++u) *reinterpret_cast<uint4*>(a.uc + (i + u * kArThreads) * 16) = t[u];
    }
    for (; i < hi; i += kArThreads)
      *reinterpret_cast<uint4*>(a.uc + i * 16) = Wire<false>::load_scaled(buf, i, scale);
  }
  if (b == 0 && tid < numel - tail0)
    reinterpret_cast<float*>(a.uc + tail0 * 4)[tid] = __fmul_rn(buf[tail0 + tid], scale);
  __threadfence_system();
  if (!cross_barrier<W>(v, epoch, timeout_ns)) { poison_local(v, numel); return; }

  // in-switch reduce + broadcast of my slice
  {
    const long long slo = min((long long)rank * Q, npack), shi = min(slo + Q, npack);
    const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
    long long i = lo + tid;
    for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
      uint4 t[U];
#pragma unroll
      for (int u = 0; u < U; ++u) t[u] = multimem_ld_reduce_f32x4(a.mc + (i + u * kArThreads) * 16);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        sq += sq4(t[u]);
        multimem_st_f32x4(a.mc + (i + u * kArThreads) * 16, t[u]);
      }
    }
    for (; i < hi; i += kArThreads) {
      const uint4 t = multimem_ld_reduce_f32x4(a.mc + i * 16);
      sq += sq4(t);
      multimem_st_f32x4(a.mc + i * 16, t);
    }
    if (rank == 0 && b == 0 && tid < numel - tail0) {
      char* p = a.mc + tail0 * 4 + (size_t)tid * 4;
      const float t = multimem_ld_reduce_f32(p);
      sq = __fmaf_rn(t, t, sq);
      multimem_st_f32(p, t);
    }
  }
  if (v.sumsq) block_sumsq<W>(v, sq, true);
  __threadfence_system();
  if (!cross_barrier<W>(v, epoch + 1, timeout_ns)) { poison_local(v, numel); return; }

  // copy-out (system-scope loads: peers wrote this memory through the switch)
#pragma unroll 1
  for (int s = 0; s < W; ++s) {
    const long long slo = min((long long)s * Q, npack), shi = min(slo + Q, npack);
    const long long lo = min(slo + (long long)b * q, shi), hi = min(lo + q, shi);
    long long i = lo + tid;
    for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
      uint4 t[U];
#pragma unroll
      for (int u = 0; u < U; ++u) t[u] = ld_peer16(a.uc + (i + u * kArThreads) * 16);
#pragma unroll
      for (int u = 0; u < U; ++u) *reinterpret_cast<uint4*>(buf + (i + u * kArThreads) * 4) = t[u];
    }
    for (; i < hi; i += kArThreads) *reinterpret_cast<uint4*>(buf + i * 4) = ld_peer16(a.uc + i * 16);
  }
  if (b == 0 && tid < numel - tail0)
    buf[tail0 + tid] = Wire<false>::tail_load(a.uc + tail0 * 4, tid);
  if (v.sumsq) fold_sumsq(v, 0, W);
}

}  // namespace b2

using b2::Drv;

extern "C" int b2_nvls_create(b2_comm* c, size_t bytes, int* fd_out, size_t* rounded_out) {
  using namespace b2;
  B2_REQUIRE(c && fd_out && rounded_out, "nvls_create: null pointer");
  B2_REQUIRE(c->rank == 0, "nvls_create: only rank 0 creates the multicast object");
  B2_REQUIRE(c->world >= 2, "nvls_create: needs at least two ranks");
  Drv* d;
  int rc = drv(&d);
  if (rc != B2_OK) return rc;
  CUmulticastObjectProp prop;
  std::memset(&prop, 0, sizeof(prop));
  prop.numDevices = (unsigned)c->world;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  prop.size = bytes;
  size_t gran = 0;
  B2_DRV_CHECK(d->multicastGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  if (gran == 0) return fail(B2_ERR_CUDA, "nvls_create: zero multicast granularity");
  prop.size = (bytes + gran - 1) / gran * gran;
  B2_DRV_CHECK(d->multicastCreate(&c->nvls.mc, &prop));
  int fd = -1;
  B2_DRV_CHECK(d->memExportToShareableHandle(&fd, c->nvls.mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  c->nvls.bytes = prop.size;
  c->nvls.have_mc = true;
  *fd_out = fd;
  *rounded_out = prop.size;
  return B2_OK;
}

extern "C" int b2_nvls_import(b2_comm* c, int fd, size_t rounded_bytes) {
  using namespace b2;
  B2_REQUIRE(c && fd >= 0, "nvls_import: bad arguments");
  B2_REQUIRE(c->rank != 0, "nvls_import: rank 0 already owns the object");
  Drv* d;
  int rc = drv(&d);
  if (rc != B2_OK) return rc;
  B2_DRV_CHECK(d->memImportFromShareableHandle(&c->nvls.mc, (void*)(uintptr_t)fd,
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  c->nvls.bytes = rounded_bytes;
  c->nvls.have_mc = true;
  return B2_OK;
}

extern "C" int b2_nvls_add_device(b2_comm* c) {
  using namespace b2;
  B2_REQUIRE(c && c->nvls.have_mc, "nvls_add_device: no multicast object yet");
  Drv* d;
  int rc = drv(&d);
  if (rc != B2_OK) return rc;
  CUdevice dev;
  B2_DRV_CHECK(d->deviceGet(&dev, c->device));
  B2_DRV_CHECK(d->multicastAddDevice(c->nvls.mc, dev));
  return B2_OK;
}

extern "C" int b2_nvls_bind(b2_comm* c) {
  using namespace b2;
  B2_REQUIRE(c && c->nvls.have_mc, "nvls_bind: no multicast object yet");
  Drv* d;
  int rc = drv(&d);
  if (rc != B2_OK) return rc;
  b2_nvls& n = c->nvls;
  CUmemAllocationProp prop;
  std::memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = c->device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  B2_DRV_CHECK(d->memGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (gran == 0 || n.bytes % gran != 0)
    return fail(B2_ERR_CUDA, "nvls_bind: size %zu is not a multiple of the allocation granularity %zu", n.bytes, gran);
  B2_DRV_CHECK(d->memCreate(&n.mem, n.bytes, &prop, 0));
  B2_DRV_CHECK(d->multicastBindMem(n.mc, 0, n.mem, 0, n.bytes, 0));
  CUmemAccessDesc acc;
  std::memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  B2_DRV_CHECK(d->memAddressReserve(&n.uc_va, n.bytes, gran, 0, 0));
  B2_DRV_CHECK(d->memMap(n.uc_va, n.bytes, 0, n.mem, 0));
  B2_DRV_CHECK(d->memSetAccess(n.uc_va, n.bytes, &acc, 1));
  B2_DRV_CHECK(d->memAddressReserve(&n.mc_va, n.bytes, gran, 0, 0));
  B2_DRV_CHECK(d->memMap(n.mc_va, n.bytes, 0, n.mc, 0));
  B2_DRV_CHECK(d->memSetAccess(n.mc_va, n.bytes, &acc, 1));
  B2_CUDA_CHECK(cudaMemset((void*)n.uc_va, 0, n.bytes));
  {
    const size_t two_mib = 2ull << 20;
    size_t st = n.bytes / 2 < (64ull << 20) ? n.bytes / 2 : (64ull << 20);
    n.stage_bytes = st / two_mib * two_mib;
    if (n.stage_bytes == 0) n.stage_bytes = n.bytes;  // tiny region: staging only, no pool
  }
  n.ready = true;
  return B2_OK;
}

extern "C" int b2_allreduce_nvls(b2_comm* c, void* buf, int64_t numel, float scale, b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(c, "allreduce_nvls: null comm");
  if (!c->nvls.ready) return fail(B2_ERR_STATE, "allreduce_nvls: b2_nvls_bind has not completed on this rank");
  B2_REQUIRE(numel >= 0, "allreduce_nvls: negative numel");
  if (numel == 0) return B2_OK;
  B2_REQUIRE(buf && (((uintptr_t)buf) & 15) == 0, "allreduce_nvls: buffer null or not 16-byte aligned");
  const long long cap = (long long)((c->nvls.stage_bytes - 64) / 4 / 4 * 4);
  for (long long off = 0; off < numel; off += cap) {
    const long long n = (numel - off < cap) ? (numel - off) : cap;
    NvlsArgs a;
    fill_view_common(c, a.v);
    a.v.buf = (float*)buf + off;
    for (int r = 0; r < B2_MAX_WORLD; ++r) a.v.stage[r] = nullptr;
    if (c->sumsq) {
      a.v.sumsq = c->world;
      c->sumsq_numel += n;
    }
    a.uc = (char*)c->nvls.uc_va;
    a.mc = (char*)c->nvls.mc_va;
    const void* fn = c->world == 2 ? (const void*)allreduce_nvls_kernel<2>
                   : c->world == 4 ? (const void*)allreduce_nvls_kernel<4>
                                   : (const void*)allreduce_nvls_kernel<8>;
    // the two local copy passes are HBM-bound: give them every row of the flag pad
    const int nb = pick_blocks(n / 4 / 2, kArMaxBlocks);
    long long n_arg = n;
    float scale_arg = scale;
    uint32_t epoch = c->epoch;
    unsigned long long to = c->timeout_ns;
    void* kargs[] = {&a, &n_arg, &scale_arg, &epoch, &to};
    B2_CUDA_CHECK(cudaLaunchKernel(fn, dim3(nb), dim3(kArThreads), kargs, 0, (cudaStream_t)stream));
    c->epoch += 2;
  }
  return B2_OK;
}

// called by b2_comm_destroy
int b2_nvls_release(b2_comm* c) {
  using namespace b2;
  b2_nvls& n = c->nvls;
  if (!n.have_mc) return B2_OK;
  if (g_pool_comm == c && g_pool_next > n.stage_bytes) {
    // torch's caching allocator still owns segments inside the region (DDP buckets): unmapping
    // would leave it with dangling blocks. Keep the mappings; process teardown reclaims them.
    // The communicator object itself is about to be deleted: detach it from the pool entry points
    // and forget its bookkeeping (a later communicator starts from an empty region of its own).
    std::lock_guard<std::mutex> lock(g_pool_mutex);
    g_pool_comm = nullptr;
    g_pool_next = 0;
    g_pool_free.clear();
    return B2_OK;
  }
  Drv* d;
  if (drv(&d) != B2_OK) return B2_OK;
  if (n.ready) {
    CUdevice dev;
    d->deviceGet(&dev, c->device);
    d->memUnmap(n.mc_va, n.bytes);
    d->memAddressFree(n.mc_va, n.bytes);
    d->multicastUnbind(n.mc, dev, 0, n.bytes);
    d->memUnmap(n.uc_va, n.bytes);
    d->memAddressFree(n.uc_va, n.bytes);
    d->memRelease(n.mem);
  }
  d->memRelease(n.mc);
  n = b2_nvls();
  if (g_pool_comm == c) {
    g_pool_comm = nullptr;
    g_pool_next = 0;
    g_pool_free.clear();
  }
  return B2_OK;
}

// ---- zero-copy NVLS: the buckets themselves live in the multicast-bound region ------------------
//
// The staging variant above pays a local copy in and out, which costs what the switch saves.
// If the DDP buckets are ALLOCATED inside the multicast-bound region (a torch MemPool whose
// allocator is b2_pool_alloc below, active only while the Reducer rebuilds its buckets), every
// rank's bucket i sits at the same offset of its region, and the all-reduce is one pass over
// the slice this rank owns: v = multimem.ld_reduce [mc + off + i]; multimem.st [mc + off + i],
// v * scale - no copy, no peer reads, ~S bytes per NVLink direction in total.

namespace b2 {

__device__ __forceinline__ uint4 scale4(uint4 t, float scale) {
  t.x = __float_as_uint(__fmul_rn(__uint_as_float(t.x), scale));
  t.y = __float_as_uint(__fmul_rn(__uint_as_float(t.y), scale));
  t.z = __float_as_uint(__fmul_rn(__uint_as_float(t.z), scale));
  t.w = __float_as_uint(__fmul_rn(__uint_as_float(t.w), scale));
  return t;
}

// U switch reads (16 bytes each) in flight per thread before the first multicast store.
template <int U>
__device__ __forceinline__ void nvls_range(char* mc, long long lo, long long hi, float scale, float& sq) {
  long long i = lo + threadIdx.x;
  for (; i + (U - 1) * kArThreads < hi; i += U * kArThreads) {
    uint4 t[U];
#pragma unroll
    for (int u = 0; u < U; ++u) t[u] = multimem_ld_reduce_f32x4(mc + (i + u * kArThreads) * 16);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      t[u] = scale4(t[u], scale);
      sq += sq4(t[u]);
      multimem_st_f32x4(mc + (i + u * kArThreads) * 16, t[u]);
    }
  }
  for (; i < hi; i += kArThreads) {
    const uint4 t = scale4(multimem_ld_reduce_f32x4(mc + i * 16), scale);
    sq += sq4(t);
    multimem_st_f32x4(mc + i * 16, t);
  }
}

template <int W>
__global__ void __launch_bounds__(kArThreads, 1)
allreduce_nvls_inplace_kernel(const __grid_constant__ NvlsArgs a, long long numel, float scale,
                              uint32_t epoch, unsigned long long timeout_ns, int gated, int unroll) {
  const RankView& v = a.v;
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
  // switch reads in flight per thread: 2 by default - at world 8 the in-switch reduction is the
  // bottleneck and deeper queues only add contention (sweep in profiles/r2_allreduce_tune_w8.jsonl:
  // 128 MiB 319 us with 2, 336 us with 4, 357 us with 8); B2_AR_UNROLL = 1 -> 8, 3 -> 4
  if (unroll == 1) nvls_range<8>(a.mc, lo, hi, scale, sq);
  else if (unroll == 3) nvls_range<4>(a.mc, lo, hi, scale, sq);
  else nvls_range<2>(a.mc, lo, hi, scale, sq);
  if (rank == 0 && b == 0 && tid < numel - tail0) {
    char* p = a.mc + tail0 * 4 + (size_t)tid * 4;
    const float t = __fmul_rn(multimem_ld_reduce_f32(p), scale);
    sq = __fmaf_rn(t, t, sq);
    multimem_st_f32(p, t);
  }
  if (v.sumsq) block_sumsq<W>(v, sq, true);
  __threadfence_system();
  if (!cross_barrier<W>(v, epoch + 1, timeout_ns)) { poison_local(v, numel); return; }
  if (v.sumsq) fold_sumsq(v, 0, W);
}

}  // namespace b2

// torch.cuda.memory.CUDAPluggableAllocator entry points (alloc_fn / free_fn signatures): a bump
// allocator over the multicast-bound region of the communicator attached with
// b2_nvls_pool_attach. torch's caching allocator asks for whole segments and sub-allocates.
// Freed segments go to a sorted, coalescing free list and are handed out again first-fit, so a
// re-wrapped DDP model (a second Reducer, a second set of buckets) reuses the region instead of
// leaking it; segments are 2 MiB-granular like the caching allocator's own.
extern "C" void* b2_pool_alloc(ssize_t size, int device, void* stream) {
  (void)stream;
  std::lock_guard<std::mutex> lock(b2::g_pool_mutex);
  b2_comm* c = b2::g_pool_comm;
  if (!c || !c->nvls.ready || device != c->device || size <= 0) return nullptr;
  const size_t align = 2ull << 20;
  const size_t need = ((size_t)size + align - 1) / align * align;
  for (size_t i = 0; i < b2::g_pool_free.size(); ++i) {
    auto& f = b2::g_pool_free[i];
    if (f.second >= need) {
      const size_t off = f.first;
      if (f.second == need) b2::g_pool_free.erase(b2::g_pool_free.begin() + i);
      else { f.first += need; f.second -= need; }
      return (void*)(c->nvls.uc_va + off);
    }
  }
  if (b2::g_pool_next < c->nvls.stage_bytes) b2::g_pool_next = c->nvls.stage_bytes;
  const size_t off = (b2::g_pool_next + align - 1) / align * align;
  if (off + need > c->nvls.bytes) return nullptr;  // torch reports out-of-memory
  b2::g_pool_next = off + need;
  return (void*)(c->nvls.uc_va + off);
}

extern "C" void b2_pool_free(void* ptr, ssize_t size, int device, void* stream) {
  (void)device; (void)stream;
  std::lock_guard<std::mutex> lock(b2::g_pool_mutex);
  b2_comm* c = b2::g_pool_comm;
  if (!c || !ptr || size <= 0) return;
  const size_t align = 2ull << 20;
  const CUdeviceptr addr = (CUdeviceptr)(uintptr_t)ptr;
  // a segment of an earlier communicator's region (torch releases a MemPool lazily): not ours
  if (addr < c->nvls.uc_va || addr >= c->nvls.uc_va + c->nvls.bytes) return;
  const size_t off = (size_t)(addr - c->nvls.uc_va);
  const size_t len = ((size_t)size + align - 1) / align * align;
  auto& fl = b2::g_pool_free;
  size_t i = 0;
  while (i < fl.size() && fl[i].first < off) ++i;
  fl.insert(fl.begin() + i, std::make_pair(off, len));
  if (i + 1 < fl.size() && fl[i].first + fl[i].second == fl[i + 1].first) {  // merge with the next block
    fl[i].second += fl[i + 1].second;
    fl.erase(fl.begin() + i + 1);
  }
  if (i > 0 && fl[i - 1].first + fl[i - 1].second == fl[i].first) {          // and with the previous one
    fl[i - 1].second += fl[i].second;
    fl.erase(fl.begin() + i);
  }
  if (!fl.empty() && fl.back().first + fl.back().second == b2::g_pool_next) {  // top of the bump area
    b2::g_pool_next = fl.back().first;
    fl.pop_back();
  }
}

extern "C" int b2_nvls_pool_attach(b2_comm* c) {
  using namespace b2;
  B2_REQUIRE(c, "nvls_pool_attach: null comm");
  if (!c->nvls.ready) return fail(B2_ERR_STATE, "nvls_pool_attach: b2_nvls_bind has not completed");
  if (g_pool_comm && g_pool_comm != c) return fail(B2_ERR_STATE, "nvls_pool_attach: another communicator already backs the pool");
  g_pool_comm = c;
  return B2_OK;
}

extern "C" int b2_nvls_region(b2_comm* c, uint64_t* base_out, uint64_t* bytes_out, uint64_t* pool_offset_out) {
  using namespace b2;
  B2_REQUIRE(c && base_out && bytes_out && pool_offset_out, "nvls_region: null pointer");
  if (!c->nvls.ready) return fail(B2_ERR_STATE, "nvls_region: b2_nvls_bind has not completed");
  *base_out = (uint64_t)c->nvls.uc_va;
  *bytes_out = (uint64_t)c->nvls.bytes;
  *pool_offset_out = (uint64_t)c->nvls.stage_bytes;  // the pool hands out [pool_offset, bytes)
  return B2_OK;
}

extern "C" int b2_allreduce_nvls_inplace(b2_comm* c, uint64_t offset_bytes, int64_t numel, float scale,
                                         b2_stream_t stream) {
  using namespace b2;
  B2_REQUIRE(c, "allreduce_nvls_inplace: null comm");
  if (!c->nvls.ready) return fail(B2_ERR_STATE, "allreduce_nvls_inplace: b2_nvls_bind has not completed");
  B2_REQUIRE(numel >= 0 && (offset_bytes & 15) == 0 && offset_bytes + (uint64_t)numel * 4 <= c->nvls.bytes,
             "allreduce_nvls_inplace: [%llu, +%lld floats) is outside the %zu-byte region or misaligned",
             (unsigned long long)offset_bytes, (long long)numel, c->nvls.bytes);
  if (numel == 0) return B2_OK;
  NvlsArgs a;
  fill_view_common(c, a.v);
  a.v.buf = (float*)(c->nvls.uc_va + offset_bytes);
  for (int r = 0; r < B2_MAX_WORLD; ++r) a.v.stage[r] = nullptr;
  if (c->sumsq) {
    a.v.sumsq = c->world;
    c->sumsq_numel += numel;
  }
  a.uc = (char*)(c->nvls.uc_va + offset_bytes);
  a.mc = (char*)(c->nvls.mc_va + offset_bytes);
  uint32_t epoch = c->epoch;
  unsigned long long to = c->timeout_ns;
  int gated = 1;
  {
    ArArgs<1> g;
    g.v[0] = a.v;
    const void* gate = b2_gate_kernel_for(c->world);
    void* gargs[] = {&g, &epoch, &to};
    B2_CUDA_CHECK(cudaLaunchKernel(gate, dim3(1), dim3(32), gargs, 0, (cudaStream_t)stream));
  }
  const void* fn = c->world == 2 ? (const void*)allreduce_nvls_inplace_kernel<2>
                 : c->world == 4 ? (const void*)allreduce_nvls_inplace_kernel<4>
                                 : (const void*)allreduce_nvls_inplace_kernel<8>;
  const int nb = pick_blocks(numel / 4 / c->world, c->max_blocks, c->ppt);
  long long n_arg = numel;
  float scale_arg = scale;
  int unroll = c->unroll;
  void* kargs[] = {&a, &n_arg, &scale_arg, &epoch, &to, &gated, &unroll};
  B2_CUDA_CHECK(cudaLaunchKernel(fn, dim3(nb), dim3(kArThreads), kargs, 0, (cudaStream_t)stream));
  c->epoch += 2;
  return B2_OK;
}
