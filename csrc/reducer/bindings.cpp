// pybind surface of the native host runtime (module name: distributeddataparallel_b200._C).
#include "reducer.h"
#include "step_timer.h"
#include "vmm_arena.h"

namespace py = pybind11;
using b200ddp::Reducer;

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "B200 data-parallel engine: native reducer, bucket assignment, parameter verification";
  m.attr("DEFAULT_FIRST_BUCKET_BYTES") = b200ddp::kDefaultFirstBucketBytes;
  m.attr("DEFAULT_BUCKET_BYTES_CAP") = b200ddp::kDefaultBucketBytesCap;

  m.def("compute_bucket_assignment_by_size", &b200ddp::compute_bucket_assignment_by_size, py::arg("tensors"),
        py::arg("size_limits"), py::arg("tensor_indices") = std::vector<int64_t>{},
        py::call_guard<py::gil_scoped_release>());
  m.def("pack_param_metadata", &b200ddp::pack_param_metadata);
  m.def("compare_param_metadata", &b200ddp::compare_param_metadata);
  m.def("plan_coalesced_groups", &b200ddp::plan_coalesced_groups);

  using release = py::call_guard<py::gil_scoped_release>;
  py::class_<Reducer, std::shared_ptr<Reducer>>(m, "Reducer")
      .def(py::init<std::vector<at::Tensor>, std::vector<std::vector<int64_t>>, py::object, py::object, int64_t,
                    bool, bool, std::vector<std::string>, int64_t, int64_t>(),
           py::arg("params"), py::arg("bucket_indices"), py::arg("comm"), py::arg("allocator"),
           py::arg("world_size"), py::arg("gradient_as_bucket_view") = true,
           py::arg("find_unused_parameters") = false, py::arg("param_names") = std::vector<std::string>{},
           py::arg("first_bucket_bytes") = b200ddp::kDefaultFirstBucketBytes,
           py::arg("bucket_bytes_cap") = b200ddp::kDefaultBucketBytesCap)
      .def("prepare_for_forward", &Reducer::prepare_for_forward, release())
      .def("prepare_for_backward", &Reducer::prepare_for_backward, release())
      .def("should_rebuild_buckets", &Reducer::should_rebuild_buckets, release())
      .def("rebuild_order", &Reducer::rebuild_order)
      .def("rebuild_buckets", &Reducer::rebuild_buckets, release())
      .def("set_require_sync", &Reducer::set_require_sync)
      .def("require_sync", &Reducer::require_sync)
      .def("set_comm", &Reducer::set_comm)
      .def("set_backward_end_callback", &Reducer::set_backward_end_callback)
      .def("detach_hooks", &Reducer::detach_hooks)
      .def("attach_hooks", &Reducer::attach_hooks)
      .def("num_buckets", &Reducer::num_buckets)
      .def("bucket_sizes", &Reducer::bucket_sizes)
      .def("set_tail_splits", &Reducer::set_tail_splits)
      .def("launch_unit_sizes", &Reducer::launch_unit_sizes)
      .def("num_logical_buckets", &Reducer::num_logical_buckets)
      .def("bucket_params", &Reducer::bucket_params)
      .def("bucket_flat", &Reducer::bucket_flat)
      .def("grad_slot", &Reducer::grad_slot)
      .def("logging_data", &Reducer::logging_data)
      .def("iterations", &Reducer::iterations)
      .def("has_rebuilt", &Reducer::has_rebuilt)
      .def("note_graph_replay", &Reducer::note_graph_replay)
      .def("_autograd_hook", &Reducer::autograd_hook, release())
      .def("shadow_backward", &Reducer::shadow_backward, release())
      .def("_finalize_backward", &Reducer::finalize_backward, release());

  // Device-timed step statistics (csrc/runtime/step_timer.h): the logger's timer, fed with %globaltimer stamps.
  using b200ddp::StepSummary;
  using b200ddp::StepTimer;
  py::class_<StepTimer, std::shared_ptr<StepTimer>>(m, "StepTimer")
      .def(py::init<int64_t, int64_t>(), py::arg("skip_first") = 10, py::arg("sample_every") = 1)
      .def_static("summarize",
                  [](const std::vector<int64_t>& marks, const std::vector<int64_t>& stamps,
                     const std::vector<int64_t>& bytes, int world) {
                    const StepSummary s = StepTimer::summarize(marks, stamps, bytes);
                    py::dict d;
                    py::list units;
                    for (const auto& u : s.units) {
                      py::dict e;
                      e["unit"] = u.unit;
                      e["bytes"] = u.bytes;
                      e["us"] = u.us();
                      e["bus_gbs"] = u.bus_gbs(world);
                      units.append(e);
                    }
                    d["units"] = units;
                    if (s.has_marks) {
                      d["forward_us"] = s.forward_us;
                      d["backward_us"] = s.backward_us;
                      d["step_us"] = s.step_us;
                    }
                    if (s.has_comm) {
                      d["comm_span_us"] = s.comm_span_us;
                      d["comm_busy_us"] = s.comm_busy_us;
                      d["exposed_comm_us"] = s.exposed_comm_us;
                      d["exposed_comm_pct"] = s.exposed_comm_pct;
                    }
                    return d;
                  },
                  py::arg("marks"), py::arg("stamps"), py::arg("bytes"), py::arg("world"))
      .def("record",
           [](StepTimer& t, const std::vector<int64_t>& marks, const std::vector<int64_t>& stamps,
              const std::vector<int64_t>& bytes) { return t.record(StepTimer::summarize(marks, stamps, bytes)); })
      .def("averages", &StepTimer::averages)
      .def("samples", &StepTimer::samples);

  // Symmetric arena on the CUDA VMM API + NVLink multicast (csrc/runtime/vmm_arena.h).  Stages return "" or a message.
  using b200ddp::VmmArena;
  py::class_<VmmArena, std::shared_ptr<VmmArena>>(m, "VmmArena")
      .def(py::init<int, size_t, int, int, std::string>(), py::arg("device"), py::arg("nbytes"), py::arg("rank"),
           py::arg("world"), py::arg("session"))
      .def("alloc_local", &VmmArena::alloc_local, release())
      .def("exchange", &VmmArena::exchange, release())
      .def("map_peers", &VmmArena::map_peers, release())
      .def("multicast_create", &VmmArena::multicast_create, release())
      .def("multicast_bind", &VmmArena::multicast_bind, release())
      .def("close", &VmmArena::close, release())
      .def("multicast_supported", &VmmArena::multicast_supported)
      .def("base", &VmmArena::base)
      .def("peer_base", &VmmArena::peer_base)
      .def("multicast_base", &VmmArena::multicast_base)
      .def("nbytes", &VmmArena::nbytes)
      .def("granularity", &VmmArena::granularity);
}
