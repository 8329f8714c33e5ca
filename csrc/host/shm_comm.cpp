// Shared-memory all-reduce among the processes of one host (CPU tensors).
//
// Reference counterpart: `ShareMemCommBuilder` (`deepspeed/comm/torch.py:29-35,166-171`,
// `torch.ops.deepspeed.inference_all_reduce_`) — absent from the reference tree; semantics re-implemented here.
//
// Layout of the segment `/dev/shm/<name>`:
//   Header  { magic, world, slot_bytes, barrier state }   (one cache line per counter)
//   slot[r] { slot_bytes }                                 r = 0 .. world-1
// One call moves the buffer through the slots in pieces of at most `slot_bytes`:
//   1. every rank copies its piece into its own slot                        -> barrier
//   2. rank r sums elements [r*n/world, (r+1)*n/world) over all slots in fp32 and writes the result
//      back into EVERY slot (so phase 3 reads only local-slot memory)        -> barrier
//   3. every rank copies its slot back into the caller's buffer
// The barrier is sense-reversing on a single counter; waiting spins with `pause` and yields after a while so
// oversubscribed hosts (more ranks than cores, as in CI) still make progress.
#include <atomic>
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fcntl.h>
#include <immintrin.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#define SXBH_API extern "C" __attribute__((visibility("default")))

namespace {

constexpr uint32_t kMagic = 0x53584253u;   // "SXBS"
enum { kF32 = 0, kBF16 = 1, kF16 = 2 };

struct alignas(64) Header {
  std::atomic<uint32_t> magic;
  int32_t world;
  int64_t slot_bytes;
  alignas(64) std::atomic<int32_t> arrived;
  alignas(64) std::atomic<int32_t> sense;
  alignas(64) std::atomic<int32_t> attached;
};

struct State {
  Header* hdr = nullptr;
  uint8_t* slots = nullptr;
  size_t map_bytes = 0;
  int world = 0, rank = -1;
  int64_t slot_bytes = 0;
  int local_sense = 0;
  char name[128] = {0};
} g;

inline void cpu_relax(int& spins) {
  _mm_pause();
  if (++spins > 2000) { sched_yield(); spins = 0; }
}

void barrier() {
  g.local_sense ^= 1;
  if (g.hdr->arrived.fetch_add(1, std::memory_order_acq_rel) == g.world - 1) {
    g.hdr->arrived.store(0, std::memory_order_relaxed);
    g.hdr->sense.store(g.local_sense, std::memory_order_release);
  } else {
    int spins = 0;
    while (g.hdr->sense.load(std::memory_order_acquire) != g.local_sense) cpu_relax(spins);
  }
}

inline float bf16_to_f32(uint16_t v) {
  uint32_t u = (uint32_t)v << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);   // NaN stays NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                             // round to nearest even
  return (uint16_t)(u >> 16);
}
inline float f16_to_f32(uint16_t h) { return _cvtsh_ss(h); }
inline uint16_t f32_to_f16(float f) { return _cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT); }

// Sum elements [lo, hi) over all slots, write the result into every slot.
template <typename Load, typename Store, typename T>
void reduce_slice(int64_t lo, int64_t hi, Load load, Store store) {
  constexpr int64_t kBlock = 1024;
  float acc[kBlock];
  for (int64_t b = lo; b < hi; b += kBlock) {
    const int64_t n = (hi - b < kBlock) ? hi - b : kBlock;
    const T* s0 = reinterpret_cast<const T*>(g.slots) + b;
#pragma omp simd
    for (int64_t i = 0; i < n; ++i) acc[i] = load(s0[i]);
    for (int r = 1; r < g.world; ++r) {
      const T* s = reinterpret_cast<const T*>(g.slots + (size_t)r * g.slot_bytes) + b;
#pragma omp simd
      for (int64_t i = 0; i < n; ++i) acc[i] += load(s[i]);
    }
    for (int r = 0; r < g.world; ++r) {
      T* d = reinterpret_cast<T*>(g.slots + (size_t)r * g.slot_bytes) + b;
#pragma omp simd
      for (int64_t i = 0; i < n; ++i) d[i] = store(acc[i]);
    }
  }
}

}  // namespace

// Create (rank 0) or attach to the segment.  Returns 0, or -errno.
SXBH_API int sxbh_shm_init(const char* name, int world, int rank, long long slot_bytes) {
  if (g.hdr) return 0;
  if (world < 1 || rank < 0 || rank >= world || slot_bytes < 4096 || (slot_bytes & 63)) return -EINVAL;
  snprintf(g.name, sizeof(g.name), "/%s", name);
  const size_t bytes = sizeof(Header) + (size_t)world * (size_t)slot_bytes;
  int fd = -1;
  if (rank == 0) {
    shm_unlink(g.name);
    fd = shm_open(g.name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return -errno;
    if (ftruncate(fd, (off_t)bytes) != 0) { int e = errno; close(fd); shm_unlink(g.name); return -e; }
  } else {
    for (int tries = 0; tries < 30000; ++tries) {          // up to ~30 s for rank 0 to create it
      fd = shm_open(g.name, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && (size_t)st.st_size >= bytes) break;
        close(fd);
        fd = -1;
      }
      usleep(1000);
    }
    if (fd < 0) return -ETIMEDOUT;
  }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) return -errno;
  g.hdr = reinterpret_cast<Header*>(p);
  g.slots = reinterpret_cast<uint8_t*>(p) + sizeof(Header);
  g.map_bytes = bytes;
  g.world = world; g.rank = rank; g.slot_bytes = slot_bytes; g.local_sense = 0;
  if (rank == 0) {
    g.hdr->world = world;
    g.hdr->slot_bytes = slot_bytes;
    g.hdr->arrived.store(0);
    g.hdr->sense.store(0);
    g.hdr->attached.store(0);
    g.hdr->magic.store(kMagic, std::memory_order_release);
  } else {
    int spins = 0;
    while (g.hdr->magic.load(std::memory_order_acquire) != kMagic) cpu_relax(spins);
    if (g.hdr->world != world || g.hdr->slot_bytes != slot_bytes) return -EPROTO;
  }
  g.hdr->attached.fetch_add(1);
  int spins = 0;
  while (g.hdr->attached.load(std::memory_order_acquire) < world) cpu_relax(spins);
  return 0;
}

// In-place sum over all ranks of the host.  dtype: 0 fp32, 1 bf16, 2 fp16 (accumulation always in fp32).
SXBH_API int sxbh_shm_allreduce(void* buf, long long numel, int dtype) {
  if (!g.hdr) return -ENOTCONN;
  if (numel <= 0) return 0;
  const int es = dtype == kF32 ? 4 : 2;
  if (dtype != kF32 && dtype != kBF16 && dtype != kF16) return -EINVAL;
  const int64_t per_piece = g.slot_bytes / es;
  uint8_t* mine = g.slots + (size_t)g.rank * g.slot_bytes;
  for (int64_t off = 0; off < numel; off += per_piece) {
    const int64_t n = (numel - off < per_piece) ? numel - off : per_piece;
    memcpy(mine, reinterpret_cast<uint8_t*>(buf) + off * es, (size_t)n * es);
    barrier();
    const int64_t lo = n * g.rank / g.world, hi = n * (g.rank + 1) / g.world;
    if (dtype == kF32)
      reduce_slice<float (*)(float), float (*)(float), float>(lo, hi, [](float v) { return v; }, [](float v) { return v; });
    else if (dtype == kBF16)
      reduce_slice<float (*)(uint16_t), uint16_t (*)(float), uint16_t>(lo, hi, bf16_to_f32, f32_to_bf16);
    else
      reduce_slice<float (*)(uint16_t), uint16_t (*)(float), uint16_t>(lo, hi, f16_to_f32, f32_to_f16);
    barrier();
    memcpy(reinterpret_cast<uint8_t*>(buf) + off * es, mine, (size_t)n * es);
    barrier();        // nobody may start filling its slot for the next piece while others still read
  }
  return 0;
}

SXBH_API int sxbh_shm_barrier(void) {
  if (!g.hdr) return -ENOTCONN;
  barrier();
  return 0;
}

SXBH_API int sxbh_shm_finalize(void) {
  if (!g.hdr) return 0;
  barrier();
  munmap(g.hdr, g.map_bytes);
  if (g.rank == 0) shm_unlink(g.name);
  g = State();
  return 0;
}
