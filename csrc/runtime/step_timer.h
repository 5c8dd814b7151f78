// Device-timed step statistics (host side, C++).
//
// Role in the reference stack: the DDP wrapper created at /root/reference/dpp.py:39 owns a c10d::Logger / Timer
// (SURVEY.md §2.2 C11, T/include/torch/csrc/distributed/c10d/logger.hpp:47-127, reducer.hpp:33,166-170) that reports
// average forward / backward-compute / backward-comm / overlap times, sampled after the first iterations.  The stock
// timer records CUDA events from the host as the step is issued -- which measures nothing once the step is a replayed
// CUDA graph.  Here the time stamps are written ON THE DEVICE: the collective kernels store %globaltimer at entry and
// exit into a per-launch-unit slot (csrc/kernels/comm.cu), four 1-thread kernels mark forward start / backward start /
// backward end / step end on the training stream (b200_stamp).  This class turns one set of stamps into the figures
// the logger reports -- including the communication left exposed after the backward pass -- and keeps running
// averages with the stock sampling cadence.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace b200ddp {

struct UnitTime {
  int64_t unit = 0, bytes = 0, start_ns = 0, end_ns = 0;
  double us() const { return (end_ns - start_ns) / 1e3; }
  // bus bandwidth in GB/s with the 2(N-1)/N convention of the benchmark metric
  double bus_gbs(int world) const {
    const double f = world > 1 ? 2.0 * (world - 1) / world : 0.0;
    return end_ns > start_ns ? bytes * f / static_cast<double>(end_ns - start_ns) : 0.0;
  }
};

struct StepSummary {
  bool has_marks = false;
  double forward_us = 0, backward_us = 0, step_us = 0;
  bool has_comm = false;
  double comm_span_us = 0, comm_busy_us = 0, exposed_comm_us = 0, exposed_comm_pct = 0;
  std::vector<UnitTime> units;       // the reductions that belong to this step
};

class StepTimer {
 public:
  // marks: {forward start, backward start, backward end, step end} in ns of %globaltimer (0 = not stamped);
  // stamps: (start, end) per launch unit; bytes: payload per launch unit.
  static StepSummary summarize(const std::vector<int64_t>& marks, const std::vector<int64_t>& stamps,
                               const std::vector<int64_t>& bytes);

  // Fold one step into the running averages.  Like the stock logger, the first `skip_first` steps are ignored and
  // afterwards every `sample_every`-th call is kept (1 = all).
  explicit StepTimer(int64_t skip_first = 10, int64_t sample_every = 1) : skip_first_(skip_first), sample_every_(sample_every) {}
  bool record(const StepSummary& s);
  std::map<std::string, double> averages() const;
  int64_t samples() const { return samples_; }

 private:
  int64_t skip_first_, sample_every_, seen_ = 0, samples_ = 0;
  double fwd_ = 0, bwd_ = 0, step_ = 0, span_ = 0, busy_ = 0, exposed_ = 0;
};

}  // namespace b200ddp
