// pybind11 surface of the C++ reducer + autograd integration (post-accumulate-grad hooks
// registered from C++, so the per-parameter hook path never enters the Python interpreter).
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/csrc/autograd/function_hook.h>
#include <torch/csrc/autograd/variable.h>
#include <torch/extension.h>

#include <memory>

#include "reducer.hpp"

namespace py = pybind11;
using dptr::Reducer;

namespace {

// Fires after AccumulateGrad stored the gradient of one parameter.  If autograd produced the
// gradient outside the arena slice (generic modules), it is moved into the slice with one
// device copy and `.grad` is re-pointed at the slice; then the reducer is told.
struct ArenaGradHook : public torch::autograd::PostAccumulateGradHook {
    ArenaGradHook(std::shared_ptr<Reducer> r, int idx, at::Tensor slot)
        : reducer(std::move(r)), index(idx), grad_slot(std::move(slot)) {}
    void operator()(const torch::autograd::Variable& v) override {
        at::Tensor g = v.grad();
        if (g.defined() && g.data_ptr() != grad_slot.data_ptr()) {
            at::NoGradGuard ng;
            grad_slot.copy_(g, /*non_blocking=*/true);
            v.mutable_grad() = grad_slot;
        }
        const auto stream = at::cuda::getCurrentCUDAStream();
        reducer->mark_ready(index, reinterpret_cast<uintptr_t>(stream.stream()));
    }
    std::shared_ptr<Reducer> reducer;
    int index;
    at::Tensor grad_slot;
};

void register_hooks(std::shared_ptr<Reducer> r, std::vector<at::Tensor> params,
                    std::vector<at::Tensor> slots) {
    TORCH_CHECK(params.size() == slots.size(), "params / slots size mismatch");
    for (size_t i = 0; i < params.size(); ++i) {
        torch::autograd::impl::set_post_acc_grad_hooks(
            params[i], std::make_unique<ArenaGradHook>(r, static_cast<int>(i), slots[i]));
    }
}

uintptr_t current_stream() {
    return reinterpret_cast<uintptr_t>(at::cuda::getCurrentCUDAStream().stream());
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.doc() = "dptb200 C++ runtime: gradient reducer, bucket planner";
    m.def("compute_bucket_assignment", &dptr::compute_bucket_assignment, py::arg("nbytes"),
          py::arg("dtype_key"), py::arg("limits"));
    m.def("current_stream", &current_stream);
    py::class_<dptr::BucketSpec>(m, "BucketSpec")
        .def(py::init<>())
        .def_readwrite("offset", &dptr::BucketSpec::offset)
        .def_readwrite("numel", &dptr::BucketSpec::numel)
        .def_readwrite("params", &dptr::BucketSpec::params);
    py::class_<dptr::SgdFusion>(m, "SgdFusion")
        .def(py::init<>())
        .def_readwrite("enabled", &dptr::SgdFusion::enabled)
        .def_readwrite("param_ptrs", &dptr::SgdFusion::param_ptrs)
        .def_readwrite("shadow_ptrs", &dptr::SgdFusion::shadow_ptrs)
        .def_readwrite("momentum", &dptr::SgdFusion::momentum)
        .def_readwrite("lr", &dptr::SgdFusion::lr)
        .def_readwrite("mu", &dptr::SgdFusion::mu)
        .def_readwrite("wd", &dptr::SgdFusion::wd)
        .def_readwrite("nesterov", &dptr::SgdFusion::nesterov)
        .def_readwrite("param_mc", &dptr::SgdFusion::param_mc)
        .def_readwrite("shadow_mc", &dptr::SgdFusion::shadow_mc);
    py::class_<dptr::BucketTiming>(m, "BucketTiming")
        .def_readonly("ready_ms", &dptr::BucketTiming::ready_ms)
        .def_readonly("start_ms", &dptr::BucketTiming::start_ms)
        .def_readonly("end_ms", &dptr::BucketTiming::end_ms)
        .def_readonly("bytes", &dptr::BucketTiming::bytes);
    py::class_<Reducer, std::shared_ptr<Reducer>>(m, "Reducer")
        .def(py::init<std::vector<dptr::BucketSpec>, int, uintptr_t, int, std::vector<uintptr_t>,
                      std::vector<uintptr_t>, uintptr_t, int, bool>())
        .def("set_fusion", &Reducer::set_fusion)
        .def("set_lr", &Reducer::set_lr)
        .def("set_scale_options", &Reducer::set_scale_options)
        .def("set_enabled", &Reducer::set_enabled)
        .def("set_rank", &Reducer::set_rank)
        .def("set_nvls", &Reducer::set_nvls)
        .def("set_mc_push", &Reducer::set_mc_push)
        .def("set_oneshot", &Reducer::set_oneshot)
        .def("set_local_only", &Reducer::set_local_only)
        .def("set_relaxed_barriers", &Reducer::set_relaxed_barriers)
        .def("bucket_algos", &Reducer::bucket_algos)
        .def("prepare_for_backward", &Reducer::prepare_for_backward)
        .def("mark_ready", &Reducer::mark_ready)
        .def("finalize", &Reducer::finalize)
        .def("all_launched", &Reducer::all_launched)
        .def("timings", &Reducer::timings)
        .def("backward_ms", &Reducer::backward_ms)
        .def("launches", &Reducer::launches)
        .def("comm_stream", &Reducer::comm_stream)
        .def("buckets", &Reducer::buckets);
    m.def("register_hooks", &register_hooks);
}
