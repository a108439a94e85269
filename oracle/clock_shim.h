/* Frame clock for the reference build (test infrastructure only).
 *
 * The reference's frequency tracker gates on std::chrono::steady_clock
 * (src/dsp/fft_process.cpp:349-361), which is not reproducible offline.  This
 * header is force-included (-include) in front of the UNMODIFIED reference
 * translation units; it pulls in every standard header they use first and then
 * renames `steady_clock` to a clock whose now() is set by the harness to
 * t = (frameIndex+1) * N / fs  (note T).  No reference code is
 * changed; only its time source. */
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <complex>
#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <deque>
#include <functional>
#include <mutex>
#include <numeric>
#include <queue>
#include <string>
#include <thread>
#include <vector>

namespace std { namespace chrono {
struct sdr_frame_clock {
    typedef nanoseconds duration;
    typedef duration::rep rep;
    typedef duration::period period;
    typedef chrono::time_point<sdr_frame_clock, duration> time_point;
    static constexpr bool is_steady = true;
    static time_point now() noexcept;   /* defined in ref_harness.cpp */
};
} }
#define steady_clock sdr_frame_clock
