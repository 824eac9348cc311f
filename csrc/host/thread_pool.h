// Fixed-size worker pool with a blocking parallel_for; the host-side counterpart of the reference's Engine.default /
// Engine.model pools (S/utils/ThreadPool.scala:35-230: invokeAndWait over a fixed executor).
#pragma once
#include <condition_variable>
#include <cstddef>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace bdlh {

class ThreadPool {
 public:
  static ThreadPool& instance();
  int size() const { return static_cast<int>(workers_.size()) + 1; }
  void resize(int threads);
  // Runs fn(begin, end) over [0, n) split into at most size() contiguous chunks of at least `grain` items; the calling
  // thread takes one chunk. Returns when all chunks are done. Exceptions in fn terminate (fn must not throw).
  void parallel_for(size_t n, size_t grain, const std::function<void(size_t, size_t)>& fn);
  ~ThreadPool();

 private:
  ThreadPool();
  void start(int threads);
  void stop();
  void worker_loop();

  std::vector<std::thread> workers_;
  std::mutex mu_;
  std::condition_variable cv_work_, cv_done_;
  std::vector<std::function<void()>> queue_;
  size_t pending_ = 0;
  bool stopping_ = false;
  std::mutex run_mu_;   // one parallel_for at a time
};

}  // namespace bdlh
