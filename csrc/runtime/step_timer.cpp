// See step_timer.h.
#include "step_timer.h"

#include <algorithm>

namespace b200ddp {

StepSummary StepTimer::summarize(const std::vector<int64_t>& marks, const std::vector<int64_t>& stamps,
                                 const std::vector<int64_t>& bytes) {
  StepSummary s;
  const int64_t f0 = marks.size() > 0 ? marks[0] : 0, b0 = marks.size() > 1 ? marks[1] : 0;
  const int64_t b1 = marks.size() > 2 ? marks[2] : 0, s1 = marks.size() > 3 ? marks[3] : 0;
  if (f0 && b0 >= f0) {
    s.has_marks = true;
    s.forward_us = (b0 - f0) / 1e3;
  }
  if (b0 && b1 >= b0) s.backward_us = (b1 - b0) / 1e3;
  if (f0 && s1 > f0) s.step_us = (s1 - f0) / 1e3;
  // reductions of THIS step: started after the backward pass began (a unit of the previous step that still holds an
  // older stamp, e.g. after a bucket rebuild, is left out) and, when the step end is known, not later than it
  int64_t first = 0, last = 0;
  for (size_t i = 0; i + 1 < stamps.size() && i / 2 < bytes.size(); i += 2) {
    UnitTime u;
    u.unit = static_cast<int64_t>(i / 2);
    u.bytes = bytes[i / 2];
    u.start_ns = stamps[i];
    u.end_ns = stamps[i + 1];
    if (!u.start_ns || u.end_ns < u.start_ns || !u.bytes) continue;
    if (b0 && u.start_ns < b0) continue;
    if (s1 > b0 && u.end_ns > s1 + 10'000'000) continue;
    s.units.push_back(u);
    first = first ? std::min(first, u.start_ns) : u.start_ns;
    last = std::max(last, u.end_ns);
    s.comm_busy_us += u.us();
  }
  if (!s.units.empty() && b1) {
    s.has_comm = true;
    s.comm_span_us = (last - first) / 1e3;
    s.exposed_comm_us = std::max<double>(0.0, (last - b1) / 1e3);
    if (s.step_us > 0) s.exposed_comm_pct = 100.0 * s.exposed_comm_us / s.step_us;
  }
  return s;
}

bool StepTimer::record(const StepSummary& s) {
  const int64_t n = seen_++;
  if (n < skip_first_ || !s.has_marks) return false;
  if (sample_every_ > 1 && (n - skip_first_) % sample_every_ != 0) return false;
  ++samples_;
  fwd_ += s.forward_us;
  bwd_ += s.backward_us;
  step_ += s.step_us;
  span_ += s.comm_span_us;
  busy_ += s.comm_busy_us;
  exposed_ += s.exposed_comm_us;
  return true;
}

std::map<std::string, double> StepTimer::averages() const {
  const double n = samples_ ? static_cast<double>(samples_) : 1.0;
  std::map<std::string, double> m;
  m["device_timed_samples"] = static_cast<double>(samples_);
  m["avg_forward_device_time_us"] = fwd_ / n;
  m["avg_backward_device_time_us"] = bwd_ / n;
  m["avg_step_device_time_us"] = step_ / n;
  m["avg_comm_span_device_time_us"] = span_ / n;
  m["avg_comm_busy_device_time_us"] = busy_ / n;
  m["avg_exposed_comm_device_time_us"] = exposed_ / n;
  m["avg_exposed_comm_pct"] = step_ > 0 ? 100.0 * exposed_ / step_ : 0.0;
  return m;
}

}  // namespace b200ddp
