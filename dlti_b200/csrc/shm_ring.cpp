// Shared-memory control ring for the tensor-parallel workers of one node (what vLLM's
// shm_broadcast MessageQueue does for its workers). The driver rank publishes one message per engine step (the
// packed step metadata, a few KB); every worker reads every message in order. One writer, up to 16 readers, a
// ring of fixed-size slots in a POSIX shared-memory object; sequence counters with acquire/release ordering, no
// locks, no system calls on the fast path (readers spin briefly, then yield). Replaces two gloo broadcasts per step.
// No reference counterpart (the reference ships no serving code).
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <thread>

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#define DLTI_API extern "C" __attribute__((visibility("default")))

namespace {

constexpr int MAX_READERS = 16;
constexpr uint64_t MAGIC = 0x646c74695f72696eULL;   // "dlti_rin"

struct alignas(64) Cursor {
  std::atomic<uint64_t> seq;
  char pad[56];
};
struct Header {
  uint64_t magic;
  uint32_t n_slots, slot_bytes, n_readers, _pad;
  Cursor written;                 // messages published so far
  Cursor read[MAX_READERS];       // messages consumed by reader r
};
struct Ring {
  Header* h;
  uint8_t* slots;
  size_t map_bytes;
  int reader;                     // -1 = writer
  char name[128];
};

inline size_t slot_stride(uint32_t slot_bytes) { return (sizeof(uint64_t) + slot_bytes + 63) & ~size_t(63); }

inline void backoff(int& spins) {
  if (++spins < 2000) {
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  } else if (spins < 20000) {
    sched_yield();
  } else {
    std::this_thread::sleep_for(std::chrono::microseconds(50));
  }
}

}  // namespace

// Create (writer, reader < 0) or attach to (reader index 0..n_readers-1) the ring `name`. Returns a handle or NULL.
DLTI_API void* dlti_shm_ring_open(const char* name, int reader, int n_readers, int n_slots, int slot_bytes) {
  if (!name || n_readers < 1 || n_readers > MAX_READERS || n_slots < 2 || slot_bytes < 8) return nullptr;
  const size_t bytes = sizeof(Header) + (size_t)n_slots * slot_stride((uint32_t)slot_bytes);
  int fd;
  if (reader < 0) {
    shm_unlink(name);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) { if (fd >= 0) close(fd); return nullptr; }
  } else {
    fd = -1;
    for (int i = 0; i < 20000 && fd < 0; ++i) {          // the writer may not have created it yet (<= 20 s)
      fd = shm_open(name, O_RDWR, 0600);
      if (fd < 0) std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    if (fd < 0) return nullptr;
    struct stat st;
    for (int i = 0; i < 20000; ++i) {                    // ... or not sized it yet
      if (fstat(fd, &st) == 0 && (size_t)st.st_size >= bytes) break;
      std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
  }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) return nullptr;
  Ring* r = new Ring();
  r->h = reinterpret_cast<Header*>(p);
  r->slots = reinterpret_cast<uint8_t*>(p) + sizeof(Header);
  r->map_bytes = bytes;
  r->reader = reader;
  std::strncpy(r->name, name, sizeof(r->name) - 1);
  if (reader < 0) {
    r->h->n_slots = (uint32_t)n_slots;
    r->h->slot_bytes = (uint32_t)slot_bytes;
    r->h->n_readers = (uint32_t)n_readers;
    r->h->written.seq.store(0, std::memory_order_relaxed);
    for (int i = 0; i < MAX_READERS; ++i) r->h->read[i].seq.store(0, std::memory_order_relaxed);
    std::atomic_thread_fence(std::memory_order_release);
    reinterpret_cast<std::atomic<uint64_t>*>(&r->h->magic)->store(MAGIC, std::memory_order_release);
  } else {
    int spins = 0;
    while (reinterpret_cast<std::atomic<uint64_t>*>(&r->h->magic)->load(std::memory_order_acquire) != MAGIC) backoff(spins);
  }
  return r;
}

// Writer: publish one message (blocks while the slowest reader is a whole ring behind). 0 on success.
DLTI_API int dlti_shm_ring_write(void* handle, const void* data, long long nbytes, int timeout_ms) {
  Ring* r = reinterpret_cast<Ring*>(handle);
  if (!r || r->reader >= 0 || nbytes < 0 || nbytes > r->h->slot_bytes) return -1;
  const uint64_t w = r->h->written.seq.load(std::memory_order_relaxed);
  const auto t0 = std::chrono::steady_clock::now();
  int spins = 0;
  for (;;) {
    uint64_t slowest = w;
    for (uint32_t i = 0; i < r->h->n_readers; ++i) {
      const uint64_t s = r->h->read[i].seq.load(std::memory_order_acquire);
      if (s < slowest) slowest = s;
    }
    if (w - slowest < r->h->n_slots) break;
    backoff(spins);
    if (timeout_ms > 0 && (spins & 1023) == 0 &&
        std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(timeout_ms)) return -2;
  }
  uint8_t* slot = r->slots + (w % r->h->n_slots) * slot_stride(r->h->slot_bytes);
  *reinterpret_cast<uint64_t*>(slot) = (uint64_t)nbytes;
  std::memcpy(slot + sizeof(uint64_t), data, (size_t)nbytes);
  r->h->written.seq.store(w + 1, std::memory_order_release);
  return 0;
}

// Reader: copy the next message into `out` (capacity bytes). Returns its size, -2 on timeout, -1 on error.
DLTI_API long long dlti_shm_ring_read(void* handle, void* out, long long capacity, int timeout_ms) {
  Ring* r = reinterpret_cast<Ring*>(handle);
  if (!r || r->reader < 0) return -1;
  Cursor& mine = r->h->read[r->reader];
  const uint64_t s = mine.seq.load(std::memory_order_relaxed);
  const auto t0 = std::chrono::steady_clock::now();
  int spins = 0;
  while (r->h->written.seq.load(std::memory_order_acquire) <= s) {
    backoff(spins);
    if (timeout_ms > 0 && (spins & 1023) == 0 &&
        std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(timeout_ms)) return -2;
  }
  const uint8_t* slot = r->slots + (s % r->h->n_slots) * slot_stride(r->h->slot_bytes);
  const uint64_t n = *reinterpret_cast<const uint64_t*>(slot);
  if ((long long)n > capacity) return -1;
  std::memcpy(out, slot + sizeof(uint64_t), (size_t)n);
  mine.seq.store(s + 1, std::memory_order_release);
  return (long long)n;
}

DLTI_API void dlti_shm_ring_close(void* handle) {
  Ring* r = reinterpret_cast<Ring*>(handle);
  if (!r) return;
  const bool writer = r->reader < 0;
  munmap(r->h, r->map_bytes);
  if (writer) shm_unlink(r->name);
  delete r;
}
