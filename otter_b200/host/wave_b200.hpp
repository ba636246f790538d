// otter_b200/host/wave_b200.hpp — software pipeline for assemble_process's loop body.
//
// The reference works one region at a time per host thread (src/assemble.cpp:43-44,103-142).  On
// the GPU hot path a region's alignments only pay off in batches, and the stages alternate between
// device and host:
//     device 1  every region's distance matrix            (fill_dist_matrices)
//     host   1  clustering per region                     (otter_hclust, host threads)
//     device 2  re-assignment + consensus alignments      (invalid_reassignments,
//                                                          rapid_consensus_alignments)
//     host   2  partial-order consensus per region        (rapid_consensus_finish, host threads)
// run_waves() cuts the regions into waves and overlaps them: while the device works on the
// distance matrices of wave i+1 the host threads cluster wave i, while the device aligns wave i
// against its representatives the host threads cluster wave i+1 and build the consensus of wave
// i-1.  Host stages get [first,last) region ranges and may themselves use many threads.
// Device stages are issued in wave order, by default from the calling thread only (one context per
// aligner, one stream).  issuers = 2 issues device 2 from a second thread with its own aligners, so
// that the host-side packing and unpacking inside one device stage overlaps the kernels of the
// other; the kernels themselves never overlap (device_turn.hpp: batched calls take the device's
// turn, first come first served).  Measured (profiles/README.md, r1m): +13-16 % on runs of ten
// C2 or C1 waves, but every hand-over between the issuers is a thread wake-up on a host whose
// cores are busy with the host stages, and with short device stages (C1 at 3-5 waves) that costs
// more than the overlap saves — hence opt-in.
// Results do not depend on the wave size or the number of issuers (every stage is per-region work).
#pragma once
#include <algorithm>
#include <cstdint>
#include <functional>
#include <future>
#include <thread>
#include <vector>

namespace otb200 {

struct WaveStages {
  std::function<void(uint32_t, uint32_t)> device1, host1, device2, host2;
};

constexpr uint32_t kMinWavesForTwoIssuers = 4;

inline void run_waves(uint32_t n_regions, uint32_t wave, const WaveStages& s, int issuers = 1) {
  if (n_regions == 0) return;
  if (wave == 0 || wave > n_regions) wave = n_regions;
  const uint32_t n_waves = (n_regions + wave - 1) / wave;
  auto lo = [&](uint32_t w) { return w * wave; };
  auto hi = [&](uint32_t w) { return std::min(n_regions, (w + 1) * wave); };
  std::vector<std::future<void>> clustered(n_waves), finished(n_waves);
  if (issuers < 2 || n_waves < kMinWavesForTwoIssuers) {
    for (uint32_t w = 0; w < n_waves; ++w) {
      s.device1(lo(w), hi(w));
      clustered[w] = std::async(std::launch::async, s.host1, lo(w), hi(w));
      if (w > 0) {
        clustered[w - 1].get();
        s.device2(lo(w - 1), hi(w - 1));
        finished[w - 1] = std::async(std::launch::async, s.host2, lo(w - 1), hi(w - 1));
      }
    }
    clustered[n_waves - 1].get();
    s.device2(lo(n_waves - 1), hi(n_waves - 1));
    s.host2(lo(n_waves - 1), hi(n_waves - 1));
    for (uint32_t w = 0; w + 1 < n_waves; ++w) finished[w].get();
    return;
  }
  // two issuers: wave w enters device 2 as soon as its clustering is done, whatever device 1 is at
  std::vector<std::promise<void>> handed(n_waves);  // clustered[w] has been set up
  std::thread second([&] {
    for (uint32_t w = 0; w < n_waves; ++w) {
      handed[w].get_future().wait();
      clustered[w].get();
      s.device2(lo(w), hi(w));
      finished[w] = std::async(std::launch::async, s.host2, lo(w), hi(w));
    }
  });
  for (uint32_t w = 0; w < n_waves; ++w) {
    s.device1(lo(w), hi(w));
    clustered[w] = std::async(std::launch::async, s.host1, lo(w), hi(w));
    handed[w].set_value();
  }
  second.join();
  for (uint32_t w = 0; w < n_waves; ++w) finished[w].get();
}

}  // namespace otb200
