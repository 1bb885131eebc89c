// A small persistent pool of host threads that drains pinned staging chunks into pageable destinations.
// ugu::Image keeps its pixels in pageable memory (a std::vector, include/ugu/image.h:354), so the Render
// boundary for C++ hosts ends in a host-side copy of 24 bytes per pixel; one thread moves about 12 GB/s, the
// PCIe link delivers 57 GB/s, so the copy is spread over a few threads, each waiting for the DMA of the chunk
// it takes (CUDA events) and then copying it.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace ugu_b200 {

struct HostChunk {
  void* dst;
  const void* stage;
  size_t bytes;
  cudaEvent_t ready;  // recorded behind the chunk's device -> staging DMA
};

class HostCopyPool {
 public:
  HostCopyPool(int device, int workers) : device_(device) {
    for (int i = 0; i < workers; i++) threads_.emplace_back([this] { Loop(); });
  }
  ~HostCopyPool() {
    {
      std::lock_guard<std::mutex> l(m_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  int workers() const { return (int)threads_.size(); }

  // Copies every chunk (the calling thread takes part); returns the first CUDA error seen.
  cudaError_t Run(const std::vector<HostChunk>& chunks) {
    {
      std::lock_guard<std::mutex> l(m_);
      chunks_ = &chunks;
      next_.store(0);
      error_.store((int)cudaSuccess);
      active_ = (int)threads_.size();
      generation_++;
    }
    cv_.notify_all();
    Drain();
    std::unique_lock<std::mutex> l(m_);
    done_.wait(l, [this] { return active_ == 0; });
    chunks_ = nullptr;
    return (cudaError_t)error_.load();
  }

 private:
  void Drain() {
    const std::vector<HostChunk>& c = *chunks_;
    for (;;) {
      const size_t i = next_.fetch_add(1);
      if (i >= c.size()) return;
      const cudaError_t e = cudaEventSynchronize(c[i].ready);
      if (e != cudaSuccess) {
        error_.store((int)e);
        continue;
      }
      memcpy(c[i].dst, c[i].stage, c[i].bytes);
    }
  }
  void Loop() {
    cudaSetDevice(device_);
    unsigned long long seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> l(m_);
        cv_.wait(l, [&] { return stop_ || generation_ != seen; });
        if (stop_) return;
        seen = generation_;
      }
      Drain();
      {
        std::lock_guard<std::mutex> l(m_);
        active_--;
      }
      done_.notify_one();
    }
  }

  int device_;
  std::vector<std::thread> threads_;
  std::mutex m_;
  std::condition_variable cv_, done_;
  const std::vector<HostChunk>* chunks_ = nullptr;
  std::atomic<size_t> next_{0};
  std::atomic<int> error_{0};
  int active_ = 0;
  unsigned long long generation_ = 0;
  bool stop_ = false;
};

}  // namespace ugu_b200
