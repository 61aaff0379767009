// C++ autograd hooks of the gradient reducer (a small torch extension; the only file of the framework that includes
// torch headers).
//
// torch DDP gives the reference (/root/reference/multigpu.py:89) c10d::Reducer, whose per-parameter hooks are C++
// post-hooks on the gradient accumulators (c10d/reducer.hpp:219-224, autograd_hook :74) - no Python runs per gradient.
// This is the same mechanism: a LambdaPostHook per parameter on its AccumulateGrad node that hands the gradient's device
// pointer to the native reducer (csrc/runtime/reducer.cu, through a plain function pointer - this module does not link
// against libddpx_kernels.so and contains no CUDA).  Python is entered twice per iteration at most: when the first
// bucket has been launched (the early buffer broadcast) and when the last one has (finalisation, queued on the engine
// exactly like c10d's finalize_backward, reducer.hpp:286).
#include <torch/extension.h>
#include <torch/csrc/autograd/engine.h>
#include <torch/csrc/autograd/utils/lambda_post_hook.h>

#include <atomic>
#include <memory>
#include <vector>

namespace {

using MarkFn = int (*)(void* handle, int pid, const void* grad_ptr, void* stream);

class HookSet {
 public:
  HookSet(std::vector<at::Tensor> params, py::object on_first_launch, py::object on_all_launched)
      : params_(std::move(params)), on_first_(std::move(on_first_launch)), on_done_(std::move(on_all_launched)) {
    state_ = std::make_shared<State>();
    for (size_t i = 0; i < params_.size(); ++i) {
      auto acc = torch::autograd::impl::grad_accumulator(params_[i]);
      TORCH_CHECK(acc, "parameter ", i, " has no gradient accumulator (does it require a gradient?)");
      const int pid = static_cast<int>(i);
      auto st = state_;
      at::Tensor p = params_[i];
      HookSet* self = this;
      auto key = acc->add_post_hook(std::make_unique<torch::autograd::utils::LambdaPostHook>(
          [st, self, p, pid](const torch::autograd::variable_list& outputs,
                             const torch::autograd::variable_list& /*inputs*/) -> torch::autograd::variable_list {
            if (st->alive.load() && st->armed.load()) self->on_grad(p, pid);
            return outputs;
          }));
      accs_.push_back(acc);
      keys_.push_back(key);
    }
  }

  ~HookSet() { remove(); }

  void set_native(uintptr_t handle, uintptr_t mark_fn) {
    handle_ = reinterpret_cast<void*>(handle);
    mark_ = reinterpret_cast<MarkFn>(mark_fn);
  }

  // prepare_for_backward: hooks fire from now on; `stream` is the raw handle of the stream backward's kernels run on
  void arm(uintptr_t stream) {
    stream_ = reinterpret_cast<void*>(stream);
    queued_ = false;
    state_->armed.store(true);
  }
  void disarm() { state_->armed.store(false); }      // no_sync(): gradients just accumulate
  bool armed() const { return state_->armed.load(); }

  void remove() {
    if (!state_ || !state_->alive.exchange(false)) return;
    for (size_t i = 0; i < accs_.size(); ++i) accs_[i]->del_post_hook(keys_[i]);
    accs_.clear();
    keys_.clear();
  }

 private:
  struct State {
    std::atomic<bool> alive{true};
    std::atomic<bool> armed{false};
  };

  void on_grad(const at::Tensor& p, int pid) {
    TORCH_CHECK(mark_ != nullptr && handle_ != nullptr, "ddpx reducer hooks: native reducer not attached");
    at::Tensor g = p.grad();
    TORCH_CHECK(g.defined(), "gradient hook fired without a gradient (parameter ", pid, ")");
    if (!g.is_contiguous() || g.scalar_type() != p.scalar_type()) {
      // the bucket gather reads the gradient as a dense array of the parameter's dtype
      g = g.to(p.scalar_type()).contiguous();
      p.mutable_grad() = g;
    }
    const int flags = mark_(handle_, pid, g.data_ptr(), stream_);
    if (flags == -1)
      throw std::runtime_error("parameter " + std::to_string(pid) +
                               " was marked ready twice in one iteration (used twice in the forward? reducer.hpp:559-568)");
    if (flags < 0) throw std::runtime_error("ddpx native reducer failed, code " + std::to_string(flags));
    if ((flags & 4) && !on_first_.is_none()) {
      py::gil_scoped_acquire gil;
      on_first_();
    }
    if ((flags & 2) && !queued_) {
      queued_ = true;
      state_->armed.store(false);            // everything of this iteration has been launched
      // the callback runs at the end of this backward pass; python objects are only touched with the GIL held
      HookSet* self = this;
      auto st = state_;
      torch::autograd::Engine::get_default_engine().queue_callback([self, st]() {
        if (!st->alive.load()) return;
        py::gil_scoped_acquire gil;
        self->on_done_();
      });
    }
  }

  std::vector<at::Tensor> params_;
  py::object on_first_, on_done_;
  std::shared_ptr<State> state_;
  std::vector<std::shared_ptr<torch::autograd::Node>> accs_;
  std::vector<uintptr_t> keys_;
  void* handle_ = nullptr;
  MarkFn mark_ = nullptr;
  void* stream_ = nullptr;
  bool queued_ = false;
};

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "C++ autograd post-hooks feeding the native gradient reducer";
  py::class_<HookSet>(m, "HookSet")
      .def(py::init<std::vector<at::Tensor>, py::object, py::object>(), py::arg("params"), py::arg("on_first_launch"),
           py::arg("on_all_launched"))
      .def("set_native", &HookSet::set_native)
      .def("arm", &HookSet::arm)
      .def("disarm", &HookSet::disarm)
      .def("armed", &HookSet::armed)
      .def("remove", &HookSet::remove);
}
