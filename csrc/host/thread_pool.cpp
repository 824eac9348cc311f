#include "thread_pool.h"

#include <algorithm>
#include <cstdlib>

#include "bigdl_b200_host.h"

namespace bdlh {

static int default_threads() {
  if (const char* e = std::getenv("BIGDL_B200_HOST_THREADS")) {
    const int v = std::atoi(e);
    if (v > 0) return v;
  }
  unsigned hc = std::thread::hardware_concurrency();
  if (hc == 0) hc = 4;
  return static_cast<int>(std::min(hc, 16u));   // data-loader helpers, not a compute pool
}

ThreadPool& ThreadPool::instance() {
  static ThreadPool pool;
  return pool;
}

ThreadPool::ThreadPool() { start(default_threads()); }
ThreadPool::~ThreadPool() { stop(); }

void ThreadPool::start(int threads) {
  stopping_ = false;
  for (int i = 1; i < threads; ++i) workers_.emplace_back([this] { worker_loop(); });
}

void ThreadPool::stop() {
  {
    std::lock_guard<std::mutex> g(mu_);
    stopping_ = true;
  }
  cv_work_.notify_all();
  for (auto& t : workers_) t.join();
  workers_.clear();
}

void ThreadPool::resize(int threads) {
  std::lock_guard<std::mutex> run(run_mu_);
  if (threads < 1) threads = 1;
  if (threads == size()) return;
  stop();
  start(threads);
}

void ThreadPool::worker_loop() {
  for (;;) {
    std::function<void()> job;
    {
      std::unique_lock<std::mutex> lk(mu_);
      cv_work_.wait(lk, [this] { return stopping_ || !queue_.empty(); });
      if (queue_.empty()) return;   // stopping
      job = std::move(queue_.back());
      queue_.pop_back();
    }
    job();
    {
      std::lock_guard<std::mutex> g(mu_);
      if (--pending_ == 0) cv_done_.notify_all();
    }
  }
}

void ThreadPool::parallel_for(size_t n, size_t grain, const std::function<void(size_t, size_t)>& fn) {
  if (n == 0) return;
  if (grain == 0) grain = 1;
  size_t chunks = std::min<size_t>(static_cast<size_t>(size()), (n + grain - 1) / grain);
  if (chunks <= 1) {
    fn(0, n);
    return;
  }
  std::lock_guard<std::mutex> run(run_mu_);
  const size_t per = (n + chunks - 1) / chunks;
  {
    std::lock_guard<std::mutex> g(mu_);
    for (size_t c = 1; c < chunks; ++c) {
      const size_t b = c * per, e = std::min(n, b + per);
      if (b >= e) continue;
      queue_.emplace_back([&fn, b, e] { fn(b, e); });
      ++pending_;
    }
  }
  cv_work_.notify_all();
  fn(0, std::min(n, per));
  std::unique_lock<std::mutex> lk(mu_);
  cv_done_.wait(lk, [this] { return pending_ == 0; });
}

}  // namespace bdlh

extern "C" int bdlh_pool_size(void) { return bdlh::ThreadPool::instance().size(); }
extern "C" void bdlh_pool_resize(int threads) { bdlh::ThreadPool::instance().resize(threads); }
extern "C" const char* bdlh_version(void) { return "bigdl_b200_host 0.1"; }
