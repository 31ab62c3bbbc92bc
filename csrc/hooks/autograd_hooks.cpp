// C++ autograd hooks of the DistributedDataParallel wrapper.
//
// c10d's Reducer registers its gradient hooks and its end-of-backward callback in C++, so the hot per-parameter path
// of a backward pass never takes the GIL (the reference reaches it through DistributedDataParallel,
// ref:pytorch/resnet/main.py:44-46,131; ref:pytorch/unet/train.py:68-70,191).  This module is the counterpart for the
// sm_100a data path:
//   * every trainable parameter gets a post-accumulate-grad hook object (torch::autograd::PostAccumulateGradHook)
//     that re-homes a freshly installed gradient tensor into its bucket view and marks the parameter ready in the
//     reducer core (libdmpi_host.so: dmpi_reducer_mark_ready, which launches the bucket's NVLink all-reduce through the
//     CUDA callback of libdmpi_cuda.so);
//   * the first mark of a backward pass queues the finalize step on the autograd engine (Engine::queue_callback):
//     unfinished-bucket check + "compute stream waits for the last all-reduce";
//   * the fused backward kernels' wrappers report readiness through `mark_ready`, a pybind entry that drops the GIL.
// The reducer core and the CUDA glue are reached through plain function pointers handed over at creation, so this
// library links against libtorch only and the two dmpi libraries stay free of any torch dependency.
#include <torch/csrc/autograd/engine.h>
#include <torch/csrc/autograd/function_hook.h>
#include <torch/csrc/autograd/variable.h>
#include <torch/extension.h>

#include <atomic>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

using mark_ready_fn = int (*)(void*, int, int*, int);
using pending_fn = int (*)(void*);
using finalize_fn = int (*)(void*);

struct HookCtx {
  void* reducer = nullptr;
  void* cuda_red = nullptr;  // may be null (no CUDA glue: the caller launches the returned buckets itself)
  mark_ready_fn mark = nullptr;
  pending_fn pending = nullptr;
  finalize_fn finalize = nullptr;
  int n_buckets = 0;
  std::mutex mu;
  bool armed = false;
  bool finalize_queued = false;
  std::vector<uint8_t> marked;
  std::vector<at::Tensor> views;       // bucket view of every parameter's gradient (logical shape)
  const uint8_t* fused = nullptr;      // per parameter: gradient written by a fused kernel (never published by a hook)
  at::Tensor fused_keepalive;
  std::vector<std::string> names;
  long long launched_buckets = 0;      // buckets handed to the launch callback so far (launch accounting)
  long long completed = 0;             // backward passes finalized so far
};

void finalize_backward(const std::shared_ptr<HookCtx>& c) {
  {
    std::lock_guard<std::mutex> lk(c->mu);
    c->finalize_queued = false;
    if (!c->armed) return;
    c->armed = false;
  }
  const int pend = c->pending(c->reducer);
  if (pend != 0)
    throw std::runtime_error("DDP: backward finished with " + std::to_string(pend) +
                             " gradient bucket(s) incomplete - some parameters did not receive a gradient (unused "
                             "parameters are not supported)");
  if (c->cuda_red != nullptr) {
    const int rc = c->finalize(c->cuda_red);
    if (rc != 0) throw std::runtime_error("DDP: all-reduce launch failed (" + std::to_string(rc) + ")");
  }
  std::lock_guard<std::mutex> lk(c->mu);
  ++c->completed;
}

// Marks parameter `idx`; must run on an autograd engine thread during a backward pass (queue_callback needs the
// current graph task).  Returns the number of buckets launched by this call.
int mark_ready_impl(const std::shared_ptr<HookCtx>& c, int idx) {
  bool queue = false;
  {
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->armed) return 0;
    if (idx < 0 || idx >= static_cast<int>(c->marked.size())) throw std::out_of_range("DDP: bad parameter index");
    if (!c->finalize_queued) {
      c->finalize_queued = true;
      queue = true;
    }
    c->marked[idx] = 1;
  }
  if (queue) torch::autograd::Engine::get_default_engine().queue_callback([c]() { finalize_backward(c); });
  std::vector<int> out(static_cast<size_t>(c->n_buckets > 0 ? c->n_buckets : 1));
  const int n = c->mark(c->reducer, idx, out.data(), c->n_buckets);
  if (n == -1)
    throw std::runtime_error("DDP: parameter '" + c->names[idx] +
                             "' was marked ready twice in one backward pass (shared module used twice?)");
  if (n > 0) {
    std::lock_guard<std::mutex> lk(c->mu);
    c->launched_buckets += n;
  }
  return n;
}

struct ReadyHook : public torch::autograd::PostAccumulateGradHook {
  std::shared_ptr<HookCtx> ctx;
  int idx;
  ReadyHook(std::shared_ptr<HookCtx> c, int i) : ctx(std::move(c)), idx(i) {}
  void operator()(const torch::autograd::Variable& param) override {
    HookCtx& c = *ctx;
    if (c.fused != nullptr && c.fused[idx]) return;  // published by the fused kernel's wrapper only
    const at::Tensor& view = c.views[idx];
    const at::Tensor& g = param.grad();
    if (g.defined() && g.data_ptr() != view.data_ptr()) {
      // autograd installed a fresh tensor (e.g. after zero_grad(set_to_none=True)): move it into the bucket view
      at::NoGradGuard ng;
      view.copy_(g);
      param.mutable_grad() = view;
    }
    {
      std::lock_guard<std::mutex> lk(c.mu);
      if (!c.armed || c.marked[idx]) return;
    }
    mark_ready_impl(ctx, idx);
  }
};

std::shared_ptr<HookCtx> create(uint64_t reducer, uint64_t cuda_red, uint64_t mark_fn, uint64_t pending_fn_addr,
                                uint64_t finalize_fn_addr, int n_buckets) {
  auto c = std::make_shared<HookCtx>();
  c->reducer = reinterpret_cast<void*>(reducer);
  c->cuda_red = reinterpret_cast<void*>(cuda_red);
  c->mark = reinterpret_cast<mark_ready_fn>(mark_fn);
  c->pending = reinterpret_cast<pending_fn>(pending_fn_addr);
  c->finalize = reinterpret_cast<finalize_fn>(finalize_fn_addr);
  c->n_buckets = n_buckets;
  return c;
}

void install(const std::shared_ptr<HookCtx>& c, const std::vector<at::Tensor>& params,
             const std::vector<at::Tensor>& views, const std::vector<std::string>& names, at::Tensor fused_flags) {
  TORCH_CHECK(params.size() == views.size() && params.size() == names.size(), "params / views / names differ in length");
  TORCH_CHECK(fused_flags.device().is_cpu() && fused_flags.scalar_type() == at::kByte && fused_flags.is_contiguous() &&
                  fused_flags.numel() == static_cast<int64_t>(params.size()),
              "fused_flags must be a contiguous CPU uint8 tensor with one entry per parameter");
  c->views = views;
  c->names = names;
  c->marked.assign(params.size(), 0);
  c->fused_keepalive = fused_flags;
  c->fused = fused_flags.data_ptr<uint8_t>();
  for (size_t i = 0; i < params.size(); ++i) {
    TORCH_CHECK(params[i].requires_grad(), "parameter '", names[i], "' does not require grad");
    torch::autograd::impl::set_post_acc_grad_hooks(params[i], std::make_unique<ReadyHook>(c, static_cast<int>(i)));
  }
}

void arm(const std::shared_ptr<HookCtx>& c, bool on) {
  std::lock_guard<std::mutex> lk(c->mu);
  c->armed = on;
  if (on) std::fill(c->marked.begin(), c->marked.end(), 0);
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "C++ autograd hooks of deeplearning_mpi_b200.parallel.ddp (post-accumulate-grad hooks + finalize callback)";
  py::class_<HookCtx, std::shared_ptr<HookCtx>>(m, "HookCtx");
  m.def("create", &create, "reducer, cuda_reducer, &dmpi_reducer_mark_ready, &dmpi_reducer_pending, &dmpi_reducer_cuda_finalize, n_buckets");
  m.def("install", &install);
  m.def("arm", &arm);
  m.def("mark_ready", &mark_ready_impl, py::call_guard<py::gil_scoped_release>(),
        "report parameter idx ready from a fused backward wrapper (runs without the GIL)");
  m.def("is_marked", [](const std::shared_ptr<HookCtx>& c, int idx) {
    std::lock_guard<std::mutex> lk(c->mu);
    return idx >= 0 && idx < static_cast<int>(c->marked.size()) && c->marked[idx] != 0;
  });
  m.def("completed", [](const std::shared_ptr<HookCtx>& c) {
    std::lock_guard<std::mutex> lk(c->mu);
    return c->completed;
  });
  m.def("take_launched", [](const std::shared_ptr<HookCtx>& c) {
    std::lock_guard<std::mutex> lk(c->mu);
    const long long n = c->launched_buckets;
    c->launched_buckets = 0;
    return n;
  });
}
