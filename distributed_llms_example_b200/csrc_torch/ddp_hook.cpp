// C++ DDP communication hook: takes the per-bucket Python call off the autograd thread.
//
// The Reducer (SP/torch/include/torch/csrc/distributed/c10d/reducer.hpp:97) accepts a
// CommHookInterface; torch's own default path is C++ (default_comm_hooks.hpp:36-51) while a
// user hook registered from Python costs a GIL acquisition and ~25 us of interpreter work per
// bucket, ~100 times per step, on the thread that also launches the backward kernels. This
// hook keeps a table {bucket index -> (address, numel, path, argument)} filled by the Python side
// once a bucket is registered (comm.py HookState.registration_for), and for registered buckets
// does the whole job in C++: order the side stream after the producer stream with one
// reusable event, call libb2ddp's C entry point (function pointers handed over as integers, so
// this file links nothing from libb2ddp), complete a device-aware Future with the side stream
// current. Anything not in the table (first steps, small buckets, the collective registration
// vote) is delegated to the Python hook, exactly as before.
//
// This is the only file of the package that includes torch headers; the kernels and their C
// ABI stay torch-free (include/b2ddp.h).
#include <ATen/cuda/CUDAEvent.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/csrc/distributed/c10d/comm.hpp>
#include <torch/csrc/distributed/c10d/reducer.hpp>
#include <torch/csrc/jit/python/pybind_utils.h>
#include <torch/extension.h>

#include <memory>
#include <vector>

namespace {

using RegisteredFn = int (*)(void*, int, int64_t, float, void*);       // b2_allreduce_registered
using NvlsFn = int (*)(void*, uint64_t, int64_t, float, void*);        // b2_allreduce_nvls_inplace

struct Entry {
  uintptr_t ptr = 0;
  int64_t numel = 0;
  int kind = -1;      // 0: IPC-registered bucket (arg = registration id); 1: NVLS region (arg = byte offset)
  uint64_t arg = 0;
};

struct Shared {
  std::vector<Entry> table;
  uint64_t fast_calls = 0, slow_calls = 0;
  bool closed = false;  // set by P2PComm.close() through Handle::close(): the communicator is gone
};

class B2CommHook : public c10d::CommHookInterface {
 public:
  B2CommHook(py::object hook, py::object state, uintptr_t ctx, uintptr_t fn_registered, uintptr_t fn_nvls,
             double scale, int device, int64_t side_stream_id, uintptr_t status_word,
             std::shared_ptr<Shared> shared)
      : hook_(std::move(hook)), state_(std::move(state)), ctx_(reinterpret_cast<void*>(ctx)),
        status_(reinterpret_cast<const volatile uint32_t*>(status_word)),
        fn_registered_(reinterpret_cast<RegisteredFn>(fn_registered)), fn_nvls_(reinterpret_cast<NvlsFn>(fn_nvls)),
        scale_(static_cast<float>(scale)), device_(device),
        side_(c10::cuda::CUDAStream::unpack3(side_stream_id, static_cast<c10::DeviceIndex>(device),
                                             c10::DeviceType::CUDA)),
        shared_(std::move(shared)) {}

  ~B2CommHook() override {
    py::gil_scoped_acquire gil;
    hook_ = py::object();
    state_ = py::object();
  }

  c10::intrusive_ptr<c10::ivalue::Future> runHook(c10d::GradBucket& bucket) override {
    const size_t idx = bucket.getIndex();
    at::Tensor& buf = bucket.getBufferRef();
    TORCH_CHECK(!shared_->closed, "b200 DDP hook: its communicator has been closed (P2PComm.close()); "
                "register a new hook before running backward on this DDP module");
    // one host load of the mapped status word (include/b2ddp.h, b2_comm_status_word): a peer that
    // missed a barrier poisoned the buckets with NaN; stop here, on every rank, instead of training on
    TORCH_CHECK(status_ == nullptr || *status_ == 0u, "libb2ddp error -4: a rank did not reach a cross-GPU "
                "barrier in time (B2_AR_TIMEOUT_MS); gradients since then are NaN-filled on every rank");
    if (idx < shared_->table.size()) {
      const Entry& e = shared_->table[idx];
      if (e.kind >= 0 && e.ptr == reinterpret_cast<uintptr_t>(buf.data_ptr()) && e.numel == buf.numel()) {
        auto producer = c10::cuda::getCurrentCUDAStream(static_cast<c10::DeviceIndex>(device_));
        event_.record(producer);
        event_.block(side_);
        const int rc = e.kind == 1 ? fn_nvls_(ctx_, e.arg, e.numel, scale_, side_.stream())
                                   : fn_registered_(ctx_, static_cast<int>(e.arg), e.numel, scale_, side_.stream());
        TORCH_CHECK(rc == 0, "libb2ddp all-reduce failed with status ", rc, " on DDP bucket ", idx);
        auto fut = c10::make_intrusive<c10::ivalue::Future>(c10::TensorType::get(),
                                                             std::vector<c10::Device>{buf.device()});
        {
          c10::cuda::CUDAStreamGuard guard(side_);
          fut->markCompleted(c10::IValue(buf));  // records the completion event on the side stream
        }
        ++shared_->fast_calls;
        return fut;
      }
    }
    ++shared_->slow_calls;
    py::gil_scoped_acquire gil;
    py::object py_fut = hook_(state_, bucket);
    return py_fut.cast<std::shared_ptr<torch::jit::PythonFutureWrapper>>()->fut;
  }

  at::Tensor parseHookResult(const c10::IValue& result) override {
    if (result.isTensor()) return result.toTensor();
    TORCH_INTERNAL_ASSERT(result.isPyObject(), "expected the Python hook's future to hold a Python object");
    py::gil_scoped_acquire gil;
    py::object obj = torch::jit::toPyObject(result);
    return torch::jit::toIValue(obj, c10::TensorType::get()).toTensor();
  }

 private:
  py::object hook_, state_;
  void* ctx_;
  const volatile uint32_t* status_;
  RegisteredFn fn_registered_;
  NvlsFn fn_nvls_;
  float scale_;
  int device_;
  c10::cuda::CUDAStream side_;
  at::cuda::CUDAEvent event_;
  std::shared_ptr<Shared> shared_;
};

// Python-visible handle to the table shared with the hook owned by the Reducer.
struct Handle {
  std::shared_ptr<Shared> shared = std::make_shared<Shared>();
  void set_entry(size_t index, uintptr_t ptr, int64_t numel, int kind, uint64_t arg) {
    if (shared->table.size() <= index) shared->table.resize(index + 1);
    shared->table[index] = Entry{ptr, numel, kind, arg};
  }
  void close() {
    shared->closed = true;
    shared->table.clear();
  }
  uint64_t fast_calls() const { return shared->fast_calls; }
  uint64_t slow_calls() const { return shared->slow_calls; }
};

std::shared_ptr<Handle> attach(py::object reducer, py::object hook, py::object state, uintptr_t ctx,
                               uintptr_t fn_registered, uintptr_t fn_nvls, double scale, int device,
                               int64_t side_stream_id, uintptr_t status_word) {
  auto r = reducer.cast<std::shared_ptr<c10d::Reducer>>();
  auto handle = std::make_shared<Handle>();
  r->register_comm_hook(std::make_unique<B2CommHook>(std::move(hook), std::move(state), ctx, fn_registered,
                                                      fn_nvls, scale, device, side_stream_id, status_word,
                                                      handle->shared));
  return handle;
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  py::class_<Handle, std::shared_ptr<Handle>>(m, "Handle")
      .def("set_entry", &Handle::set_entry)
      .def("close", &Handle::close)
      .def_property_readonly("fast_calls", &Handle::fast_calls)
      .def_property_readonly("slow_calls", &Handle::slow_calls);
  m.def("attach", &attach, "Register the C++ comm hook on a c10d::Reducer");
}
