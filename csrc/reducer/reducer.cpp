// See reducer.h for the design notes and the reference citations.
#include "reducer.h"

#include <torch/csrc/autograd/engine.h>
#include <torch/csrc/autograd/utils/lambda_post_hook.h>
#include <torch/csrc/autograd/variable.h>

#include <algorithm>
#include <deque>
#include <map>
#include <sstream>
#include <unordered_set>

namespace b200ddp {

namespace {

constexpr int64_t kSlotAlignBytes = 256;   // TMA / 128-bit vector friendly slot starts

struct GroupKey {
  int8_t dtype;
  int8_t dev_type;
  int16_t dev_index;
  bool operator<(const GroupKey& o) const {
    return std::tie(dtype, dev_type, dev_index) < std::tie(o.dtype, o.dev_type, o.dev_index);
  }
};

GroupKey key_of(const at::Tensor& t) {
  return GroupKey{static_cast<int8_t>(t.scalar_type()), static_cast<int8_t>(t.device().type()),
                  static_cast<int16_t>(t.device().index())};
}

double ms_between(std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
  return std::chrono::duration<double, std::milli>(b - a).count();
}

}  // namespace

// --------------------------------------------------------------------------------------
// bucket assignment
// --------------------------------------------------------------------------------------
std::pair<std::vector<std::vector<int64_t>>, std::vector<int64_t>> compute_bucket_assignment_by_size(
    const std::vector<at::Tensor>& tensors, const std::vector<int64_t>& size_limits,
    const std::vector<int64_t>& tensor_indices) {
  TORCH_CHECK(!size_limits.empty(), "need at least one bucket size limit");
  TORCH_CHECK(tensor_indices.empty() || tensor_indices.size() == tensors.size(),
              "tensor_indices must be empty or match tensors");
  struct Open {
    std::vector<int64_t> members;
    int64_t bytes = 0;
    size_t limit_pos = 0;   // which entry of size_limits is active for this group
  };
  struct Closed {
    std::vector<int64_t> members;
    int64_t limit;
    int64_t smallest;
  };
  std::map<GroupKey, Open> open;
  std::vector<GroupKey> first_seen;     // stable order for the leftovers
  std::vector<Closed> closed;

  auto close = [&](Open& o) {
    Closed c;
    c.limit = size_limits[o.limit_pos];
    c.smallest = *std::min_element(o.members.begin(), o.members.end());
    c.members = std::move(o.members);
    closed.push_back(std::move(c));
    o.members.clear();
    o.bytes = 0;
  };

  for (size_t i = 0; i < tensors.size(); ++i) {
    const auto& t = tensors[i];
    TORCH_CHECK(!t.is_sparse(), "sparse gradients are not supported");
    const int64_t id = tensor_indices.empty() ? static_cast<int64_t>(i) : tensor_indices[i];
    const GroupKey k = key_of(t);
    auto it = open.find(k);
    if (it == open.end()) {
      it = open.emplace(k, Open{}).first;
      first_seen.push_back(k);
    }
    Open& o = it->second;
    o.members.push_back(id);
    o.bytes += t.numel() * static_cast<int64_t>(t.element_size());
    if (o.bytes >= size_limits[o.limit_pos]) {
      close(o);
      if (o.limit_pos + 1 < size_limits.size()) ++o.limit_pos;
    }
  }
  for (const auto& k : first_seen) {
    Open& o = open[k];
    if (!o.members.empty()) close(o);
  }
  if (tensor_indices.empty()) {
    std::stable_sort(closed.begin(), closed.end(),
                     [](const Closed& a, const Closed& b) { return a.smallest < b.smallest; });
  }
  std::pair<std::vector<std::vector<int64_t>>, std::vector<int64_t>> out;
  for (auto& c : closed) {
    out.first.push_back(std::move(c.members));
    out.second.push_back(c.limit);
  }
  return out;
}

// --------------------------------------------------------------------------------------
// parameter metadata (constructor-time verification)
// --------------------------------------------------------------------------------------
at::Tensor pack_param_metadata(const std::vector<at::Tensor>& params) {
  std::vector<int64_t> rec;
  rec.push_back(static_cast<int64_t>(params.size()));
  for (const auto& p : params) {
    rec.push_back(static_cast<int64_t>(p.scalar_type()));
    rec.push_back(p.dim());
    for (auto s : p.sizes()) rec.push_back(s);
    for (auto s : p.strides()) rec.push_back(s);
  }
  return at::tensor(rec, at::TensorOptions().dtype(at::kLong));
}

std::string compare_param_metadata(const at::Tensor& mine, const at::Tensor& theirs, int64_t their_rank) {
  auto a = mine.contiguous().cpu();
  auto b = theirs.contiguous().cpu();
  const int64_t* pa = a.data_ptr<int64_t>();
  const int64_t* pb = b.data_ptr<int64_t>();
  std::ostringstream msg;
  if (pa[0] != pb[0]) {
    msg << "parameter count differs: this rank has " << pa[0] << ", rank " << their_rank << " has " << pb[0];
    return msg.str();
  }
  int64_t ia = 1, ib = 1;
  for (int64_t t = 0; t < pa[0]; ++t) {
    if (ia + 1 >= a.numel() || ib + 1 >= b.numel()) return "truncated metadata";
    const int64_t dta = pa[ia], nda = pa[ia + 1], dtb = pb[ib], ndb = pb[ib + 1];
    if (dta != dtb) {
      msg << "parameter " << t << ": dtype differs from rank " << their_rank;
      return msg.str();
    }
    if (nda != ndb) {
      msg << "parameter " << t << ": rank (ndim) differs from rank " << their_rank;
      return msg.str();
    }
    for (int64_t d = 0; d < nda; ++d) {
      if (pa[ia + 2 + d] != pb[ib + 2 + d]) {
        msg << "parameter " << t << ": size of dim " << d << " is " << pa[ia + 2 + d] << " here but "
            << pb[ib + 2 + d] << " on rank " << their_rank;
        return msg.str();
      }
    }
    for (int64_t d = 0; d < nda; ++d) {
      if (pa[ia + 2 + nda + d] != pb[ib + 2 + nda + d]) {
        msg << "parameter " << t << ": stride of dim " << d << " differs from rank " << their_rank;
        return msg.str();
      }
    }
    ia += 2 + 2 * nda;
    ib += 2 + 2 * ndb;
  }
  return "";
}

std::vector<std::vector<int64_t>> plan_coalesced_groups(const std::vector<at::Tensor>& tensors,
                                                        int64_t buffer_bytes) {
  // dtype groups keep first-appearance order; inside a dtype, tensors keep their order and a
  // group is cut before the tensor that would push it past buffer_bytes (a single tensor
  // larger than the buffer gets a group of its own).
  std::vector<std::vector<int64_t>> groups;
  std::map<int, std::pair<size_t, int64_t>> cur;   // dtype -> (group index, bytes so far)
  for (size_t i = 0; i < tensors.size(); ++i) {
    const auto& t = tensors[i];
    const int dt = static_cast<int>(t.scalar_type());
    const int64_t bytes = t.numel() * static_cast<int64_t>(t.element_size());
    auto it = cur.find(dt);
    if (it == cur.end() || (it->second.second + bytes > buffer_bytes && it->second.second > 0)) {
      groups.emplace_back();
      cur[dt] = {groups.size() - 1, 0};
      it = cur.find(dt);
    }
    groups[it->second.first].push_back(static_cast<int64_t>(i));
    it->second.second += bytes;
  }
  return groups;
}

// --------------------------------------------------------------------------------------
// Reducer
// --------------------------------------------------------------------------------------
Reducer::Reducer(std::vector<at::Tensor> params, std::vector<std::vector<int64_t>> bucket_indices,
                 py::object comm, py::object allocator, int64_t world_size, bool gradient_as_bucket_view,
                 bool find_unused_parameters, std::vector<std::string> param_names,
                 int64_t first_bucket_bytes, int64_t bucket_bytes_cap)
    : params_(std::move(params)),
      names_(std::move(param_names)),
      comm_(std::move(comm)),
      allocator_(std::move(allocator)),
      world_size_(world_size),
      grad_as_view_(gradient_as_bucket_view),
      find_unused_(find_unused_parameters),
      first_bucket_bytes_(first_bucket_bytes),
      bucket_bytes_cap_(bucket_bytes_cap) {
  TORCH_CHECK(!params_.empty(), "Reducer needs at least one parameter");
  if (names_.size() != params_.size()) {
    names_.resize(params_.size());
    for (size_t i = 0; i < names_.size(); ++i)
      if (names_[i].empty()) names_[i] = "param" + std::to_string(i);
  }
  slots_.resize(params_.size());
  ready_.assign(params_.size(), 0);
  build_buckets(bucket_indices);

  attach_hooks();
}

void Reducer::attach_hooks() {
  // One post hook per parameter on its gradient accumulator: fires after the gradient has
  // been accumulated into .grad for this backward pass (cf. reducer.hpp:74 autograd_hook).
  // The accumulator node records the CUDA stream that is current when it is created; call
  // this under the stream the training step runs on (see reattach_hooks()).
  detach_hooks();
  accumulators_.reserve(params_.size());
  for (size_t i = 0; i < params_.size(); ++i) {
    TORCH_CHECK(params_[i].requires_grad(), "parameter ", names_[i], " does not require grad");
    auto acc = torch::autograd::impl::grad_accumulator(params_[i]);
    const int64_t index = static_cast<int64_t>(i);
    auto key = acc->add_post_hook(std::make_unique<torch::autograd::utils::LambdaPostHook>(
        [this, index](const torch::autograd::variable_list& outputs,
                      const torch::autograd::variable_list& /*inputs*/) {
          this->autograd_hook(index);
          return outputs;
        }));
    accumulator_index_[acc.get()] = index;
    hook_keys_.push_back(key);
    accumulators_.push_back(std::move(acc));
  }
}

Reducer::~Reducer() {
  detach_hooks();
  // py::objects may be released without the GIL being held by the destroying thread
  py::gil_scoped_acquire gil;
  for (auto& b : buckets_) b.handle = py::object();
  comm_ = py::object();
  allocator_ = py::object();
  backward_end_cb_ = py::object();
}

void Reducer::detach_hooks() {
  for (size_t i = 0; i < accumulators_.size() && i < hook_keys_.size(); ++i) {
    if (accumulators_[i]) accumulators_[i]->del_post_hook(hook_keys_[i]);
  }
  hook_keys_.clear();
  accumulators_.clear();
  accumulator_index_.clear();
}

void Reducer::set_tail_splits(std::vector<int64_t> thresholds) {
  std::lock_guard<std::mutex> lock(mu_);
  std::sort(thresholds.begin(), thresholds.end(), std::greater<int64_t>());
  thresholds.erase(std::remove_if(thresholds.begin(), thresholds.end(), [](int64_t v) { return v <= 0; }), thresholds.end());
  tail_splits_ = std::move(thresholds);
}

std::vector<int64_t> Reducer::launch_unit_sizes() const {
  std::vector<int64_t> out;
  for (const auto& b : buckets_) out.push_back(b.payload_bytes);
  return out;
}

void Reducer::build_buckets(const std::vector<std::vector<int64_t>>& bucket_indices,
                            const std::vector<int64_t>& logical_of_unit) {
  py::gil_scoped_acquire gil;
  at::NoGradGuard no_grad;
  std::vector<uint8_t> seen(params_.size(), 0);
  std::vector<Bucket> fresh;
  fresh.reserve(bucket_indices.size());
  for (size_t bi = 0; bi < bucket_indices.size(); ++bi) {
    const auto& members = bucket_indices[bi];
    TORCH_CHECK(!members.empty(), "empty bucket");
    Bucket b;
    b.logical = logical_of_unit.empty() ? static_cast<int64_t>(bi) : logical_of_unit.at(bi);
    b.closes_logical = bi + 1 == bucket_indices.size() || logical_of_unit.empty() ||
                       logical_of_unit.at(bi + 1) != b.logical;
    const auto& first = params_.at(members[0]);
    const int64_t esize = static_cast<int64_t>(first.element_size());
    const int64_t align = std::max<int64_t>(1, kSlotAlignBytes / esize);
    int64_t cursor = 0;
    for (int64_t pi : members) {
      TORCH_CHECK(pi >= 0 && pi < static_cast<int64_t>(params_.size()), "bad parameter index ", pi);
      TORCH_CHECK(!seen[pi], "parameter ", names_[pi], " assigned to two buckets");
      seen[pi] = 1;
      const auto& p = params_[pi];
      TORCH_CHECK(p.scalar_type() == first.scalar_type() && p.device() == first.device(),
                  "bucket mixes dtypes or devices");
      b.params.push_back(pi);
      b.offsets.push_back(cursor);
      b.payload_bytes += p.numel() * esize;
      cursor += (p.numel() + align - 1) / align * align;
    }
    b.numel = std::max<int64_t>(cursor, align);
    py::object flat = allocator_(b.numel, py::cast(first.scalar_type()), py::cast(first.device()));
    b.flat = flat.cast<at::Tensor>();
    TORCH_CHECK(b.flat.numel() >= b.numel && b.flat.is_contiguous(), "allocator returned a bad bucket buffer");
    for (size_t k = 0; k < b.params.size(); ++k) {
      const auto& p = params_[b.params[k]];
      at::Tensor v;
      if (p.is_non_overlapping_and_dense()) {
        // as_strided offsets are relative to the STORAGE, and `flat` may itself be a view into a
        // larger allocation (the symmetric arena)
        v = b.flat.as_strided(p.sizes(), p.strides(), b.flat.storage_offset() + b.offsets[k]);
      } else {
        v = b.flat.narrow(0, b.offsets[k], p.numel()).view(p.sizes());
      }
      b.views.push_back(v);
      slots_[b.params[k]] = Slot{static_cast<int64_t>(bi), b.offsets[k], p.numel()};
    }
    fresh.push_back(std::move(b));
  }
  for (size_t i = 0; i < seen.size(); ++i) TORCH_CHECK(seen[i], "parameter ", names_[i], " is in no bucket");

  // carry live gradients over to the new layout (they may alias the old buckets)
  for (auto& b : fresh) {
    for (size_t k = 0; k < b.params.size(); ++k) {
      auto& p = params_[b.params[k]];
      at::Tensor& g = p.mutable_grad();
      if (g.defined()) {
        b.views[k].copy_(g);
        if (grad_as_view_) g = b.views[k];
      }
    }
  }
  buckets_ = std::move(fresh);
  logical_sizes_.clear();
  for (const auto& b : buckets_) {
    if (static_cast<int64_t>(logical_sizes_.size()) <= b.logical) logical_sizes_.resize(b.logical + 1, 0);
    logical_sizes_[b.logical] += b.payload_bytes;
  }
  reset_round();
}

void Reducer::reset_round() {
  for (auto& b : buckets_) {
    b.pending = static_cast<int64_t>(b.params.size());
    b.launched = false;
    b.handle = py::object();
  }
  std::fill(ready_.begin(), ready_.end(), 0);
  next_bucket_ = 0;
  callback_queued_ = false;
}

void Reducer::prepare_for_forward() {
  std::lock_guard<std::mutex> lock(mu_);
  t_forward_start_ = clock::now();
}

void Reducer::prepare_for_backward(const std::vector<at::Tensor>& outputs) {
  std::unique_lock<std::mutex> lock(mu_);
  TORCH_CHECK(!armed_ || next_bucket_ == 0 || next_bucket_ == num_buckets(),
              "Expected to have finished reduction in the prior iteration before starting a new one. "
              "Some parameters did not receive a gradient; pass find_unused_parameters=True.");
  {
    py::gil_scoped_acquire gil;
    reset_round();
  }
  armed_ = true;
  t_backward_start_ = clock::now();
  unused_.clear();
  if (find_unused_) {
    search_unused(outputs);
    auto unused = unused_;
    lock.unlock();
    for (int64_t i : unused) {
      std::lock_guard<std::mutex> l2(mu_);
      mark_variable_ready(i);
    }
  }
}

void Reducer::search_unused(const std::vector<at::Tensor>& outputs) {
  std::unordered_set<torch::autograd::Node*> seen;
  std::deque<torch::autograd::Node*> queue;
  for (const auto& o : outputs) {
    if (o.defined() && o.grad_fn()) queue.push_back(o.grad_fn().get());
  }
  while (!queue.empty()) {
    auto* fn = queue.front();
    queue.pop_front();
    if (!seen.insert(fn).second) continue;
    for (const auto& edge : fn->next_edges()) {
      if (auto* next = edge.function.get()) queue.push_back(next);
    }
  }
  for (const auto& kv : accumulator_index_) {
    if (!seen.count(kv.first)) unused_.push_back(kv.second);
  }
  std::sort(unused_.begin(), unused_.end());
}

void Reducer::autograd_hook(int64_t index) {
  std::lock_guard<std::mutex> lock(mu_);
  if (!armed_ || !require_sync_) return;   // no_sync(): grads accumulate locally in the slots
  if (!callback_queued_) {
    callback_queued_ = true;
    torch::autograd::Engine::get_default_engine().queue_callback([this] { this->finalize_backward(); });
  }
  mark_variable_ready(index);
}

void Reducer::mark_variable_ready(int64_t index) {
  TORCH_CHECK(index >= 0 && index < static_cast<int64_t>(params_.size()), "bad parameter index");
  TORCH_CHECK(!ready_[index], "Expected to mark a variable ready only once. Parameter ", names_[index],
              " (index ", index, ") has been marked as ready twice in this iteration; it is probably "
              "used more than once in forward under checkpointing, or hooks fired for two backward passes.");
  ready_[index] = 1;
  if (!recorded_order_ && !has_rebuilt_) ready_order_.push_back(index);
  const Slot& slot = slots_[index];
  Bucket& b = buckets_[slot.bucket];
  const size_t k = std::find(b.params.begin(), b.params.end(), index) - b.params.begin();
  {
    at::NoGradGuard no_grad;
    at::Tensor& grad = params_[index].mutable_grad();
    at::Tensor& view = b.views[k];
    if (grad.defined()) {
      if (grad.data_ptr() == view.data_ptr() && grad.strides() == view.strides()) {
        ++alias_hits_;
      } else {
        view.copy_(grad);
        ++copies_in_;
        if (grad_as_view_) grad = view;
      }
    } else {
      view.zero_();
    }
  }
  if (--b.pending == 0) launch_ready_buckets();
}

void Reducer::launch_ready_buckets() {
  // Buckets go out strictly in index order so that every rank issues the same sequence of
  // collectives even if gradients become ready in a different order (reducer.hpp:282).
  while (next_bucket_ < num_buckets() && buckets_[next_bucket_].pending == 0 && !buckets_[next_bucket_].launched) {
    Bucket& b = buckets_[next_bucket_];
    if (next_bucket_ == 0) t_comm_start_ = clock::now();
    const bool is_last = next_bucket_ + 1 == num_buckets();
    {
      py::gil_scoped_acquire gil;
      b.handle = comm_.attr("launch")(next_bucket_, b.flat, is_last);
    }
    b.launched = true;
    ++next_bucket_;
    if (b.closes_logical) ++num_buckets_reduced_total_;
  }
}

void Reducer::finalize_backward() {
  std::lock_guard<std::mutex> lock(mu_);
  if (!armed_) return;
  t_backward_end_ = clock::now();
  if (next_bucket_ != num_buckets()) {
    std::ostringstream missing;
    int shown = 0;
    for (size_t i = 0; i < ready_.size() && shown < 8; ++i) {
      if (!ready_[i]) {
        missing << (shown ? ", " : "") << names_[i];
        ++shown;
      }
    }
    armed_ = false;
    {
      py::gil_scoped_acquire gil;
      reset_round();
    }
    TORCH_CHECK(false, "Gradient reduction did not finish: parameters without a gradient this iteration: ",
                missing.str(), ". Pass find_unused_parameters=True if this is expected.");
  }
  {
    py::gil_scoped_acquire gil;
    if (backward_end_cb_ && !backward_end_cb_.is_none()) backward_end_cb_();
    for (auto& b : buckets_) {
      if (b.handle && !b.handle.is_none()) b.handle.attr("wait")();
      b.handle = py::object();
    }
  }
  t_comm_end_ = clock::now();
  {
    at::NoGradGuard no_grad;
    for (auto& b : buckets_) {
      for (size_t k = 0; k < b.params.size(); ++k) {
        at::Tensor& grad = params_[b.params[k]].mutable_grad();
        if (grad_as_view_) {
          if (!grad.defined() || grad.data_ptr() != b.views[k].data_ptr()) grad = b.views[k];
        } else if (grad.defined()) {
          grad.copy_(b.views[k]);
        } else {
          grad = b.views[k].clone();
        }
      }
    }
  }
  if (!recorded_order_ && !has_rebuilt_) recorded_order_ = true;
  // runtime statistics (host clock; the wrapper adds device-timed numbers)
  ++iterations_;
  if (t_forward_start_ != clock::time_point{}) {
    sum_forward_ms_ += ms_between(t_forward_start_, t_backward_start_);
    sum_backward_ms_ += ms_between(t_backward_start_, t_backward_end_);
    sum_comm_ms_ += ms_between(t_comm_start_, t_comm_end_);
    sum_overlap_ms_ += std::max(0.0, ms_between(t_comm_start_, std::min(t_backward_end_, t_comm_end_)));
    ++timed_iters_;
  }
  armed_ = false;
}

void Reducer::shadow_backward() {
  {
    std::unique_lock<std::mutex> lock(mu_);
    TORCH_CHECK(!armed_ || next_bucket_ == 0 || next_bucket_ == num_buckets(),
                "shadow_backward called while a reduction is in flight");
    {
      py::gil_scoped_acquire gil;
      reset_round();
    }
    armed_ = true;
    t_backward_start_ = clock::now();
    {
      at::NoGradGuard no_grad;
      for (auto& b : buckets_) {
        b.flat.zero_();
        b.pending = 0;
      }
    }
    std::fill(ready_.begin(), ready_.end(), 1);
    if (!recorded_order_ && !has_rebuilt_) {
      // no real backward ever ran on this rank: record the conventional reverse registration order
      // (rank 0's order is the one every rank adopts at the rebuild)
      ready_order_.clear();
      for (int64_t i = static_cast<int64_t>(params_.size()) - 1; i >= 0; --i) ready_order_.push_back(i);
    }
    launch_ready_buckets();
  }
  finalize_backward();
}

bool Reducer::should_rebuild_buckets() const {
  std::lock_guard<std::mutex> lock(mu_);
  return !has_rebuilt_ && recorded_order_ && !find_unused_ &&
         ready_order_.size() == params_.size();
}

std::vector<std::vector<int64_t>> Reducer::rebuild_buckets(const std::vector<int64_t>& order) {
  std::lock_guard<std::mutex> lock(mu_);
  TORCH_CHECK(order.size() == params_.size(), "rebuild order must list every parameter exactly once");
  std::vector<at::Tensor> ordered;
  ordered.reserve(order.size());
  for (int64_t i : order) ordered.push_back(params_.at(i));
  auto assignment = compute_bucket_assignment_by_size(ordered, {first_bucket_bytes_, bucket_bytes_cap_}, order);
  // launch units: the stock buckets, with the last one cut at the tail thresholds
  std::vector<std::vector<int64_t>> units;
  std::vector<int64_t> logical;
  for (size_t bi = 0; bi < assignment.first.size(); ++bi) {
    const auto& members = assignment.first[bi];
    if (bi + 1 != assignment.first.size() || tail_splits_.empty() || members.size() < 2) {
      units.push_back(members);
      logical.push_back(static_cast<int64_t>(bi));
      continue;
    }
    std::vector<int64_t> suffix(members.size() + 1, 0);       // bytes from member k to the end
    for (int64_t k = static_cast<int64_t>(members.size()) - 1; k >= 0; --k) {
      const auto& p = params_.at(members[k]);
      suffix[k] = suffix[k + 1] + p.numel() * static_cast<int64_t>(p.element_size());
    }
    size_t begin = 0;
    for (int64_t thr : tail_splits_) {
      size_t cut = begin;
      while (cut < members.size() && suffix[cut] > thr) ++cut;   // first member from which the rest fits under thr
      if (cut > begin && cut < members.size()) {
        units.emplace_back(members.begin() + begin, members.begin() + cut);
        logical.push_back(static_cast<int64_t>(bi));
        begin = cut;
      }
    }
    units.emplace_back(members.begin() + begin, members.end());
    logical.push_back(static_cast<int64_t>(bi));
  }
  build_buckets(units, logical);
  has_rebuilt_ = true;
  ready_order_.clear();
  return assignment.first;
}

std::vector<int64_t> Reducer::bucket_sizes() const { return logical_sizes_; }

std::vector<std::vector<int64_t>> Reducer::bucket_params() const {
  std::vector<std::vector<int64_t>> out;
  for (const auto& b : buckets_) out.push_back(b.params);
  return out;
}

at::Tensor Reducer::grad_slot(int64_t param_index) const {
  const Slot& s = slots_.at(param_index);
  const Bucket& b = buckets_.at(s.bucket);
  const size_t k = std::find(b.params.begin(), b.params.end(), param_index) - b.params.begin();
  return b.views.at(k);
}

py::dict Reducer::logging_data() const {
  // lock order is always mu_ -> GIL (hooks run without the GIL and take it only to call the
  // comm object), so the snapshot is taken with the GIL dropped.
  struct Snap {
    std::vector<int64_t> sizes;
    int64_t total = 0, nb = 0, reduced = 0, iters = 0, alias = 0, copies = 0, timed = 0;
    bool rebuilt = false;
    double f = 0, b = 0, c = 0, o = 0;
  } s;
  {
    py::gil_scoped_release release;
    std::lock_guard<std::mutex> lock(mu_);
    for (const auto& p : params_) s.total += p.numel() * static_cast<int64_t>(p.element_size());
    s.sizes = bucket_sizes();
    s.nb = num_logical_buckets();
    s.reduced = num_buckets_reduced_total_;
    s.iters = iterations_;
    s.alias = alias_hits_;
    s.copies = copies_in_;
    s.timed = timed_iters_;
    s.rebuilt = has_rebuilt_;
    s.f = sum_forward_ms_;
    s.b = sum_backward_ms_;
    s.c = sum_comm_ms_;
    s.o = sum_overlap_ms_;
  }
  py::dict d;
  d["num_parameter_tensors"] = static_cast<int64_t>(params_.size());
  d["total_parameter_size_bytes"] = s.total;
  d["bucket_sizes"] = s.sizes;
  d["bucket_cap_bytes"] = bucket_bytes_cap_;
  d["first_bucket_bytes"] = first_bucket_bytes_;
  d["has_rebuilt_buckets"] = s.rebuilt;
  d["num_buckets"] = s.nb;
  d["launch_unit_sizes"] = launch_unit_sizes();
  d["num_buckets_reduced"] = s.reduced;
  d["iteration"] = s.iters;
  d["world_size"] = world_size_;
  d["gradient_as_bucket_view"] = grad_as_view_;
  d["find_unused_parameters"] = find_unused_;
  d["grads_produced_in_place"] = s.alias;
  d["grads_copied_in"] = s.copies;
  const double n = s.timed ? static_cast<double>(s.timed) : 1.0;
  d["avg_forward_compute_time_ms"] = s.f / n;
  d["avg_backward_compute_time_ms"] = s.b / n;
  d["avg_backward_comm_time_ms"] = s.c / n;
  d["avg_backward_compute_comm_overlap_time_ms"] = s.o / n;
  return d;
}

}  // namespace b200ddp
