// timer.hpp -- GpuTimer / CpuTimer / GET_TIME_DIFF with the interface of
// 0-frameworks/cux/include/util/timer.h:16-81.  GpuTimer records a pair of events on the legacy
// default stream through the C ABI (cuxb_event_*), exactly where cux records them (timer.h:26-27).
#pragma once
#include <chrono>

#include "cux/core.hpp"

namespace cux {

class GpuTimer {
public:
  GpuTimer() { CUXB_CHECK(cuxb_event_create(&start_)); CUXB_CHECK(cuxb_event_create(&stop_)); }
  ~GpuTimer() { cuxb_event_destroy(start_); cuxb_event_destroy(stop_); }
  GpuTimer(const GpuTimer &) = delete;
  GpuTimer &operator=(const GpuTimer &) = delete;
  inline void Start() { CUXB_CHECK(cuxb_event_record(start_, nullptr)); }
  inline void Stop() { CUXB_CHECK(cuxb_event_record(stop_, nullptr)); }
  inline float MilliSeconds() {
    float ms = 0.f;
    CUXB_CHECK(cuxb_event_elapsed_ms(start_, stop_, &ms));  // waits for stop_
    return ms;
  }
protected:
  void *start_ = nullptr;
  void *stop_ = nullptr;
};

class CpuTimer {
public:
  typedef std::chrono::high_resolution_clock clock;
  inline void Start() { t0_ = clock::now(); }
  inline void Stop() { t1_ = clock::now(); }
  inline float NanoSeconds() { return (float)std::chrono::duration_cast<std::chrono::nanoseconds>(t1_ - t0_).count(); }
  inline float MicroSeconds() { return NanoSeconds() / 1e3f; }
  inline float MilliSeconds() { return NanoSeconds() / 1e6f; }
  inline float Seconds() { return NanoSeconds() / 1e9f; }
protected:
  std::chrono::time_point<clock> t0_, t1_;
};

}  // namespace cux

// timer.h:75-81: time a block of statements with the given timer, returning milliseconds.
#define GET_TIME_DIFF(timer, ...) \
  [&]() -> float {                \
    timer.Start();                \
    { __VA_ARGS__ }               \
    timer.Stop();                 \
    return timer.MilliSeconds();  \
  }();
