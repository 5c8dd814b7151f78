// Gradient reducer for the B200 data-parallel engine (host side, C++).
//
// Role in the reference stack: /root/reference/dpp.py:39 wraps the model in
// torch.nn.parallel.DistributedDataParallel, whose C++ c10d::Reducer (SURVEY.md §2.2 C6,
// T/include/torch/csrc/distributed/c10d/reducer.hpp:45-584) turns per-parameter
// "gradient ready" events into ordered bucket all-reduces overlapped with backward.
// This is a new implementation of that behaviour, shaped for the B200 data plane:
//
//   * gradients LIVE in the flat bucket (".grad" is a strided view into the bucket slot);
//     kernels that produce weight gradients write straight into the slot, so the per-step
//     copy-in / copy-out passes of the stock reducer disappear.  A
//     gradient that arrives in foreign memory is copied in once (generic modules).
//   * the 1/world_size averaging is NOT applied on copy-in; it is fused into the
//     reduction kernel (or applied by the c10d fallback backend).
//   * bucket memory comes from a caller-supplied allocator so that it can sit inside the
//     NVLink-mapped symmetric arena used by the peer-to-peer all-reduce kernels.
//   * launching a bucket's reduction is delegated to a "comm" object
//     (launch(index, flat, is_last) -> handle with .wait()), so the same reducer drives
//     the P2P/NVLS kernels on GPUs and gloo/NCCL on the baseline path.
#pragma once

#include <torch/extension.h>

#include <chrono>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

namespace b200ddp {

constexpr int64_t kDefaultFirstBucketBytes = 1 << 20;          // 1 MiB
constexpr int64_t kDefaultBucketBytesCap = 25 * (1 << 20);     // 25 MiB

// Greedy size-capped grouping; same observable result as stock
// dist._compute_bucket_assignment_by_size (reducer.hpp:592-598): a bucket is closed once
// its byte size reaches the active limit (so buckets may exceed the cap), limits advance
// per (dtype, device) key, leftovers are appended, and when no explicit index order is
// given buckets are sorted by their smallest tensor index.
std::pair<std::vector<std::vector<int64_t>>, std::vector<int64_t>> compute_bucket_assignment_by_size(
    const std::vector<at::Tensor>& tensors, const std::vector<int64_t>& size_limits,
    const std::vector<int64_t>& tensor_indices);

// Parameter metadata exchanged by the wrapper's constructor (reducer.hpp:600-605):
// [count, then per tensor: dtype, ndim, sizes..., strides...].
at::Tensor pack_param_metadata(const std::vector<at::Tensor>& params);
// Returns "" when equal, else a human readable description of the first mismatch.
std::string compare_param_metadata(const at::Tensor& mine, const at::Tensor& theirs, int64_t their_rank);

// Partition tensors into <= buffer_bytes groups per dtype, preserving order
// (coalesced broadcast planning: SURVEY C8, comm.hpp:13-17).
std::vector<std::vector<int64_t>> plan_coalesced_groups(const std::vector<at::Tensor>& tensors,
                                                        int64_t buffer_bytes);

struct Slot {
  int64_t bucket = -1;
  int64_t offset = 0;    // elements from the start of the bucket's flat buffer
  int64_t numel = 0;
};

struct Bucket {
  std::vector<int64_t> params;     // indices into Reducer::params_
  std::vector<int64_t> offsets;    // element offsets (aligned)
  std::vector<at::Tensor> views;   // strided views with each param's own geometry
  at::Tensor flat;
  int64_t numel = 0;               // padded element count of `flat`
  int64_t payload_bytes = 0;       // sum of param bytes (what the logger reports)
  int64_t logical = -1;            // index of the stock-layout bucket this launch unit belongs to
  bool closes_logical = true;      // last unit of its logical bucket
  int64_t pending = 0;
  bool launched = false;
  py::object handle;               // what comm.launch returned
};

class Reducer {
 public:
  Reducer(std::vector<at::Tensor> params, std::vector<std::vector<int64_t>> bucket_indices,
          py::object comm, py::object allocator, int64_t world_size, bool gradient_as_bucket_view,
          bool find_unused_parameters, std::vector<std::string> param_names,
          int64_t first_bucket_bytes, int64_t bucket_bytes_cap);
  ~Reducer();

  void prepare_for_forward();
  void prepare_for_backward(const std::vector<at::Tensor>& outputs);
  // True if the one-time regrouping (grad-ready order, [first, cap] limits) is due.
  bool should_rebuild_buckets() const;
  std::vector<int64_t> rebuild_order() const { return ready_order_; }
  // Apply rank 0's order; returns the new per-bucket parameter lists.
  std::vector<std::vector<int64_t>> rebuild_buckets(const std::vector<int64_t>& order);
  // Stream the LAST bucket: thresholds in bytes measured from the end of the bucket (e.g. {1 MiB, 64 KiB}).  The
  // bucket is cut where the gradients still to come fit under each threshold, and every piece is reduced as soon as
  // its own gradients are ready -- so only the final few KB (the stem's weight gradient) are left after backward ends.
  // The reported layout (bucket_sizes, logging) stays the stock one.  Takes effect at the next rebuild.
  void set_tail_splits(std::vector<int64_t> thresholds);
  std::vector<int64_t> launch_unit_sizes() const;
  void set_require_sync(bool v) { require_sync_ = v; }
  bool require_sync() const { return require_sync_; }
  void set_comm(py::object comm) { comm_ = std::move(comm); }
  // Called (with the GIL) at the end of the backward pass, before the bucket handles are waited on: the device-timed
  // logger stamps "backward compute done" there; the gap to the last bucket's end stamp is the exposed communication.
  void set_backward_end_callback(py::object cb) { backward_end_cb_ = std::move(cb); }
  void detach_hooks();
  // (Re)create the gradient-accumulator nodes under the CURRENT stream and hook them.  Needed
  // before capturing a step into a CUDA graph on a side stream: autograd pins an accumulator
  // to the stream it was created on, and a default-stream accumulator cannot be used while
  // another stream is capturing.
  void attach_hooks();

  // introspection
  int64_t num_buckets() const { return static_cast<int64_t>(buckets_.size()); }
  std::vector<int64_t> bucket_sizes() const;
  std::vector<std::vector<int64_t>> bucket_params() const;
  at::Tensor bucket_flat(int64_t i) const { return buckets_.at(i).flat; }
  at::Tensor grad_slot(int64_t param_index) const;
  py::dict logging_data() const;
  int64_t iterations() const { return iterations_; }
  bool has_rebuilt() const { return has_rebuilt_; }
  // Replay bookkeeping when a captured CUDA graph re-runs a step without the hooks firing.
  void note_graph_replay() { ++iterations_; num_buckets_reduced_total_ += num_logical_buckets(); }
  int64_t num_logical_buckets() const { return static_cast<int64_t>(logical_sizes_.size()); }

  // test hooks
  void autograd_hook(int64_t index);
  void finalize_backward();
  // Uneven inputs (DistributedDataParallel.join): a rank that has run out of data takes part in one
  // iteration's bucket reductions with all-zero gradients, issuing exactly the launches an active rank issues.
  void shadow_backward();

 private:
  void build_buckets(const std::vector<std::vector<int64_t>>& bucket_indices,
                     const std::vector<int64_t>& logical_of_unit = {});
  void mark_variable_ready(int64_t index);
  void launch_ready_buckets();
  void search_unused(const std::vector<at::Tensor>& outputs);
  void reset_round();

  std::vector<at::Tensor> params_;
  std::vector<std::string> names_;
  std::vector<Slot> slots_;
  std::vector<Bucket> buckets_;
  py::object comm_;
  py::object allocator_;
  py::object backward_end_cb_;
  int64_t world_size_;
  bool grad_as_view_;
  bool find_unused_;
  int64_t first_bucket_bytes_;
  int64_t bucket_bytes_cap_;
  std::vector<int64_t> tail_splits_;       // descending byte thresholds for streaming the last bucket
  std::vector<int64_t> logical_sizes_;     // payload bytes per stock-layout bucket

  std::vector<std::shared_ptr<torch::autograd::Node>> accumulators_;
  std::vector<uintptr_t> hook_keys_;
  std::unordered_map<torch::autograd::Node*, int64_t> accumulator_index_;

  mutable std::mutex mu_;
  bool armed_ = false;             // prepare_for_backward ran and hooks are expected
  bool require_sync_ = true;       // false inside no_sync()
  bool callback_queued_ = false;
  bool has_rebuilt_ = false;
  bool recorded_order_ = false;
  std::vector<uint8_t> ready_;
  std::vector<int64_t> ready_order_;
  std::vector<int64_t> unused_;
  int64_t next_bucket_ = 0;
  int64_t iterations_ = 0;
  int64_t num_buckets_reduced_total_ = 0;
  int64_t copies_in_ = 0;          // grads that had to be copied into their slot
  int64_t alias_hits_ = 0;         // grads produced in place

  using clock = std::chrono::steady_clock;
  clock::time_point t_forward_start_{}, t_backward_start_{}, t_backward_end_{}, t_comm_start_{}, t_comm_end_{};
  double sum_forward_ms_ = 0, sum_backward_ms_ = 0, sum_comm_ms_ = 0, sum_overlap_ms_ = 0;
  int64_t timed_iters_ = 0;
};

}  // namespace b200ddp
