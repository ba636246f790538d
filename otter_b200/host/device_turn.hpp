// otter_b200/host/device_turn.hpp — one batched call at a time per device.
//
// The alignment kernels are persistent whole-GPU kernels: two of them running at once time-slice
// badly (profiles/README.md, r1i).  A host process that issues batched calls from several threads
// (wave_b200.hpp issues the two device stages of consecutive waves from two threads) therefore
// takes the device's turn around the synchronous C-ABI call only: the packing before it and the
// unpacking after it run while the other thread's kernels own the GPU.
// Turns are served first come, first served (a ticket lock): with a plain mutex the thread that
// has just released the device re-takes it for its next call before the waiting thread wakes up,
// and the second issuer starves until the first has run out of waves (measured, r1m).
#pragma once
#include <condition_variable>
#include <cstdint>
#include <mutex>

namespace otb200 {

class DeviceTurn {  // BasicLockable: use with std::lock_guard / std::unique_lock
 public:
  void lock() {
    std::unique_lock<std::mutex> lk(m_);
    const uint64_t mine = next_++;
    cv_.wait(lk, [&] { return serving_ == mine; });
  }
  void unlock() {
    {
      std::lock_guard<std::mutex> lk(m_);
      ++serving_;
    }
    cv_.notify_all();
  }

 private:
  std::mutex m_;
  std::condition_variable cv_;
  uint64_t next_ = 0, serving_ = 0;
};

inline DeviceTurn& device_turn(int device) {
  static DeviceTurn turns[64];
  return turns[device >= 0 && device < 64 ? device : 0];
}

}  // namespace otb200
