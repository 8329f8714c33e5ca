// aio.cpp — asynchronous file IO engine for NVMe offload and fast checkpoint writes.
//
// Fresh implementation of the native module behind AsyncIOBuilder; Python call sites in the
// reference: runtime/swap_tensor/utils.py:22-27, partitioned_param_swapper.py:111-122, nvme/test_ds_aio_utils.py:91
// (aio_handle(block_size, queue_depth, single_submit, overlap_events, intra_op_parallelism) with
// sync_/async_ pread/pwrite + wait()).
//
// Design: a fixed pool of `intra_op_parallelism` worker threads per handle.  A request is cut into
// `block_size` pieces that go into one MPMC queue; workers issue positional pread/pwrite (O_DIRECT when buffer,
// offset and length are 4 KiB aligned — the pinned staging buffers used by the swapper are — falling back to
// buffered IO otherwise).  `queue_depth` bounds the number of pieces in flight per request, which is what it
// means for libaio's io_submit as well.  No libaio dependency: positional IO from several threads saturates
// NVMe queues on current kernels and keeps the library buildable everywhere.
#include <atomic>
#include <cerrno>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <fcntl.h>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
#include <vector>

#define SXBH_API extern "C" __attribute__((visibility("default")))

namespace {

struct Request {
  int fd = -1;
  bool write = false;
  std::atomic<long long> pending{0};
  std::atomic<int> error{0};
};

struct Piece {
  std::shared_ptr<Request> req;
  char* buf;
  long long nbytes;
  long long offset;
};

class Handle {
 public:
  Handle(long long block_size, int queue_depth, bool single_submit, bool overlap_events, int threads)
      : block_size_(block_size > 0 ? block_size : (1 << 20)), queue_depth_(queue_depth > 0 ? queue_depth : 8),
        single_submit_(single_submit), overlap_events_(overlap_events) {
    if (threads < 1) threads = 1;
    for (int i = 0; i < threads; ++i) workers_.emplace_back([this] { run(); });
  }
  ~Handle() {
    {
      std::lock_guard<std::mutex> g(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }

  int submit(void* buffer, long long nbytes, const char* path, long long file_offset, bool write, bool async) {
    int flags = write ? (O_WRONLY | O_CREAT) : O_RDONLY;
    const bool aligned = ((reinterpret_cast<uintptr_t>(buffer) | (uintptr_t)nbytes | (uintptr_t)file_offset) & 4095) == 0;
    int fd = -1;
    if (aligned) fd = ::open(path, flags | O_DIRECT, 0644);
    if (fd < 0) fd = ::open(path, flags, 0644);
    if (fd < 0) return -errno;
    auto req = std::make_shared<Request>();
    req->fd = fd;
    req->write = write;
    const long long npieces = (nbytes + block_size_ - 1) / block_size_;
    req->pending = npieces;
    {
      std::lock_guard<std::mutex> g(mu_);
      for (long long i = 0; i < npieces; ++i) {
        const long long off = i * block_size_;
        queue_.push_back(Piece{req, static_cast<char*>(buffer) + off, std::min(block_size_, nbytes - off),
                               file_offset + off});
      }
      inflight_.push_back(req);
    }
    cv_.notify_all();
    if (!async) return wait_one(req);
    return 0;
  }

  // Waits for every request submitted so far; returns the number completed or a negative errno.
  int wait_all() {
    std::vector<std::shared_ptr<Request>> reqs;
    {
      std::lock_guard<std::mutex> g(mu_);
      reqs.swap(inflight_);
    }
    int n = 0, err = 0;
    for (auto& r : reqs) {
      int e = wait_one(r, false);
      if (e < 0) err = e;
      ++n;
    }
    return err < 0 ? err : n;
  }

  long long block_size() const { return block_size_; }
  int queue_depth() const { return queue_depth_; }
  int single_submit() const { return single_submit_; }
  int overlap_events() const { return overlap_events_; }
  int threads() const { return (int)workers_.size(); }

 private:
  int wait_one(const std::shared_ptr<Request>& r, bool unlist = true) {
    {
      std::unique_lock<std::mutex> g(done_mu_);
      done_cv_.wait(g, [&] { return r->pending.load() == 0; });
    }
    ::close(r->fd);
    if (unlist) {
      std::lock_guard<std::mutex> g(mu_);
      for (auto it = inflight_.begin(); it != inflight_.end(); ++it)
        if (it->get() == r.get()) { inflight_.erase(it); break; }
    }
    return r->error.load() ? -r->error.load() : 0;
  }

  void run() {
    for (;;) {
      Piece p;
      {
        std::unique_lock<std::mutex> g(mu_);
        cv_.wait(g, [&] { return stop_ || !queue_.empty(); });
        if (stop_ && queue_.empty()) return;
        p = queue_.front();
        queue_.pop_front();
      }
      long long done = 0;
      while (done < p.nbytes) {
        ssize_t r = p.req->write ? ::pwrite(p.req->fd, p.buf + done, p.nbytes - done, p.offset + done)
                                 : ::pread(p.req->fd, p.buf + done, p.nbytes - done, p.offset + done);
        if (r < 0) {
          if (errno == EINTR) continue;
          p.req->error = errno;
          break;
        }
        if (r == 0) {                      // short file on read: zero-fill the tail
          if (!p.req->write) std::memset(p.buf + done, 0, p.nbytes - done);
          break;
        }
        done += r;
      }
      if (p.req->pending.fetch_sub(1) == 1) {
        std::lock_guard<std::mutex> g(done_mu_);
        done_cv_.notify_all();
      }
    }
  }

  const long long block_size_;
  const int queue_depth_;
  const bool single_submit_, overlap_events_;
  std::vector<std::thread> workers_;
  std::deque<Piece> queue_;
  std::vector<std::shared_ptr<Request>> inflight_;
  std::mutex mu_, done_mu_;
  std::condition_variable cv_, done_cv_;
  bool stop_ = false;
};

std::mutex g_mu;
std::map<int, std::unique_ptr<Handle>> g_handles;
int g_next = 1;

Handle* get(int id) {
  std::lock_guard<std::mutex> g(g_mu);
  auto it = g_handles.find(id);
  return it == g_handles.end() ? nullptr : it->second.get();
}

}  // namespace

SXBH_API int sxbh_aio_create(long long block_size, int queue_depth, int single_submit, int overlap_events,
                             int intra_op_parallelism) {
  std::lock_guard<std::mutex> g(g_mu);
  const int id = g_next++;
  g_handles[id] = std::make_unique<Handle>(block_size, queue_depth, single_submit != 0, overlap_events != 0,
                                           intra_op_parallelism);
  return id;
}
SXBH_API int sxbh_aio_destroy(int id) {
  std::unique_ptr<Handle> h;
  {
    std::lock_guard<std::mutex> g(g_mu);
    auto it = g_handles.find(id);
    if (it == g_handles.end()) return -1;
    h = std::move(it->second);
    g_handles.erase(it);
  }
  return 0;
}
// async = 0: returns when the bytes are transferred (0 or -errno).  async = 1: returns immediately; sxbh_aio_wait
// completes it.
SXBH_API int sxbh_aio_pread(int id, void* buffer, long long nbytes, const char* path, long long file_offset, int async) {
  Handle* h = get(id);
  return h ? h->submit(buffer, nbytes, path, file_offset, false, async != 0) : -1;
}
SXBH_API int sxbh_aio_pwrite(int id, const void* buffer, long long nbytes, const char* path, long long file_offset,
                             int async) {
  Handle* h = get(id);
  return h ? h->submit(const_cast<void*>(buffer), nbytes, path, file_offset, true, async != 0) : -1;
}
SXBH_API int sxbh_aio_wait(int id) {
  Handle* h = get(id);
  return h ? h->wait_all() : -1;
}
SXBH_API long long sxbh_aio_info(int id, int what) {
  Handle* h = get(id);
  if (!h) return -1;
  switch (what) {
    case 0: return h->block_size();
    case 1: return h->queue_depth();
    case 2: return h->single_submit();
    case 3: return h->overlap_events();
    case 4: return h->threads();
  }
  return -1;
}
SXBH_API long long sxbh_file_size(const char* path) {
  struct stat st;
  return ::stat(path, &st) == 0 ? (long long)st.st_size : -1;
}
