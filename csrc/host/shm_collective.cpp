// Shared-memory object collectives for the ranks of one box.
//
// Reference: harness/determined/core/_distributed.py:287-407 + ipc.py:34-245 move pickled objects between the workers of a
// trial over ZeroMQ (TCP through the chief) at every report / validate / checkpoint boundary, and per batch for the
// preemption check of the HF path.  On a single box all ranks can see one POSIX shared-memory segment, so an allgather is
// "write my slot, flip my sequence number, read the others": no sockets, no serialization through the chief, no syscalls on
// the fast path (waits spin briefly, then sleep).
//
// Segment layout (all offsets 64-byte aligned):
//   Header { magic, world, slot_bytes, ready }           one cache line
//   arrive[world], depart[world]                          one cache line each (no false sharing between ranks)
//   slot[world] = { uint64 total_bytes, uint64 chunk_bytes, payload[slot_bytes] }
// Protocol of round s (every rank counts rounds locally, they all make the same calls in the same order):
//   1. wait until depart[j] >= s-1 for all j      (everyone has finished reading round s-1: my slot is free)
//   2. write my slot; arrive[me] = s (release)
//   3. wait until arrive[j] >= s for all j (acquire); copy the slots out
//   4. depart[me] = s (release)
// Payloads larger than a slot take several rounds; the first round carries the total sizes.
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>

#include <fcntl.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#define DETB200_HOST_API extern "C" __attribute__((visibility("default")))

namespace {

constexpr uint64_t kMagic = 0x44455442323030ull;   // "DETB200"
constexpr size_t kLine = 64;

struct alignas(kLine) Header {
  std::atomic<uint64_t> magic;
  uint32_t world;
  int32_t creator_pid;          // attachers skip segments whose creator is gone (left behind by a crashed run)
  uint64_t slot_bytes;
};
struct alignas(kLine) Seq {
  std::atomic<uint32_t> v;
};
struct SlotHead {
  uint64_t total_bytes;
  uint64_t chunk_bytes;
};

struct Handle {
  std::string name;
  int rank = 0, world = 0;
  size_t slot_bytes = 0, map_bytes = 0;
  uint8_t* base = nullptr;
  uint32_t round = 0;
  bool owner = false;

  Header* header() const { return reinterpret_cast<Header*>(base); }
  Seq* arrive() const { return reinterpret_cast<Seq*>(base + sizeof(Header)); }
  Seq* depart() const { return arrive() + world; }
  size_t slot_stride() const { return (sizeof(SlotHead) + slot_bytes + kLine - 1) / kLine * kLine; }
  uint8_t* slot(int r) const { return base + sizeof(Header) + 2 * sizeof(Seq) * world + (size_t)r * slot_stride(); }
};

size_t segment_bytes(int world, size_t slot_bytes) {
  Handle h;
  h.world = world;
  h.slot_bytes = slot_bytes;
  return sizeof(Header) + 2 * sizeof(Seq) * world + (size_t)world * h.slot_stride();
}

thread_local char g_err[256] = "";
int fail(int code, const char* what) {
  std::snprintf(g_err, sizeof(g_err), "%s: %s", what, errno ? std::strerror(errno) : "failed");
  return code;
}

using Clock = std::chrono::steady_clock;

// spin, then yield, then sleep: object collectives are rare (boundaries), latency matters more than CPU for ~100 us
template <class Pred>
bool wait_until(Pred pred, double timeout_s) {
  for (int i = 0; i < 2000; ++i) {
    if (pred()) return true;
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
  const auto deadline = Clock::now() + std::chrono::duration<double>(timeout_s);
  int spins = 0;
  while (!pred()) {
    if (++spins < 200) std::this_thread::yield();
    else std::this_thread::sleep_for(std::chrono::microseconds(spins < 2000 ? 50 : 500));
    if ((spins & 63) == 0 && Clock::now() > deadline) return pred();
  }
  return true;
}

bool all_at_least(const Handle* h, Seq* seq, uint32_t target) {
  for (int j = 0; j < h->world; ++j)
    if ((int32_t)(seq[j].v.load(std::memory_order_acquire) - target) < 0) return false;
  return true;
}

}  // namespace

DETB200_HOST_API const char* detb200_shm_last_error(void) { return g_err; }

DETB200_HOST_API size_t detb200_shm_segment_bytes(int world, size_t slot_bytes) { return segment_bytes(world, slot_bytes); }

// Rank 0 creates and initialises the segment, the others attach (waiting for it to appear).  `name` is a POSIX shm name
// ("/detb200-<allocation>"); the segment is unlinked by rank 0 on close.
DETB200_HOST_API void* detb200_shm_open(const char* name, int rank, int world, size_t slot_bytes, double timeout_s) {
  errno = 0;
  if (world < 1 || rank < 0 || rank >= world || slot_bytes < 64) {
    fail(0, "shm_open: bad arguments");
    return nullptr;
  }
  const size_t bytes = segment_bytes(world, slot_bytes);
  int fd = -1;
  if (rank == 0) {
    shm_unlink(name);                       // a stale segment of a crashed run
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) {
      fail(0, "shm_open(create)");
      if (fd >= 0) close(fd);
      return nullptr;
    }
  } else {
    const auto deadline = Clock::now() + std::chrono::duration<double>(timeout_s);
    for (;;) {
      fd = shm_open(name, O_RDWR, 0600);
      struct stat st;
      if (fd >= 0 && fstat(fd, &st) == 0 && (size_t)st.st_size >= bytes) {
        // peek at the header: a live creator, or not initialised yet (rank 0 is between ftruncate and the magic store)
        void* q = mmap(nullptr, sizeof(Header), PROT_READ, MAP_SHARED, fd, 0);
        bool fresh = true;
        if (q != MAP_FAILED) {
          const Header* hd = static_cast<const Header*>(q);
          if (hd->magic.load(std::memory_order_acquire) == kMagic) {
            const pid_t pid = hd->creator_pid;
            fresh = pid > 0 && (kill(pid, 0) == 0 || errno == EPERM);
          }
          munmap(q, sizeof(Header));
        }
        if (fresh) break;
      }
      if (fd >= 0) close(fd);
      fd = -1;
      if (Clock::now() > deadline) {
        fail(0, "shm_open(attach): timed out waiting for rank 0");
        return nullptr;
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(2));
    }
  }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) {
    fail(0, "mmap");
    return nullptr;
  }
  auto* h = new Handle;
  h->name = name; h->rank = rank; h->world = world; h->slot_bytes = slot_bytes; h->map_bytes = bytes;
  h->base = static_cast<uint8_t*>(p);
  h->owner = rank == 0;
  if (rank == 0) {
    std::memset(p, 0, sizeof(Header) + 2 * sizeof(Seq) * world);
    h->header()->world = (uint32_t)world;
    h->header()->slot_bytes = slot_bytes;
    h->header()->creator_pid = (int32_t)getpid();
    h->header()->magic.store(kMagic, std::memory_order_release);
  } else {
    Header* hd = h->header();
    if (!wait_until([&] { return hd->magic.load(std::memory_order_acquire) == kMagic; }, timeout_s) || hd->world != (uint32_t)world ||
        hd->slot_bytes != slot_bytes) {
      errno = 0;
      fail(0, "shm_open(attach): segment not initialised or created with a different world / slot size");
      munmap(p, bytes);
      delete h;
      return nullptr;
    }
  }
  return h;
}

// Remove the name (rank 0, once every rank has attached): the mapping lives on until the last rank unmaps it, and close()
// then needs no cross-rank coordination.
DETB200_HOST_API void detb200_shm_unlink(void* handle) {
  auto* h = static_cast<Handle*>(handle);
  if (h != nullptr && h->owner) {
    shm_unlink(h->name.c_str());
    h->owner = false;
  }
}

DETB200_HOST_API void detb200_shm_close(void* handle) {
  auto* h = static_cast<Handle*>(handle);
  if (h == nullptr) return;
  munmap(h->base, h->map_bytes);
  if (h->owner) shm_unlink(h->name.c_str());
  delete h;
}

// One round: publish up to slot_bytes of `data` (with the total size of the whole message), collect everyone's chunk.
// out_chunk[j] / out_total[j] receive rank j's sizes; chunks are copied to out + j * slot_bytes.
static int round_trip(Handle* h, const uint8_t* data, uint64_t chunk, uint64_t total, uint8_t* out, uint64_t* out_chunk, uint64_t* out_total,
                      double timeout_s) {
  const uint32_t s = ++h->round;
  if (!wait_until([&] { return all_at_least(h, h->depart(), s - 1); }, timeout_s)) {
    errno = ETIMEDOUT;
    return fail(-2, "shm collective: a peer never finished the previous round");
  }
  uint8_t* mine = h->slot(h->rank);
  SlotHead hd{total, chunk};
  std::memcpy(mine, &hd, sizeof(hd));
  if (chunk) std::memcpy(mine + sizeof(hd), data, chunk);
  h->arrive()[h->rank].v.store(s, std::memory_order_release);
  if (!wait_until([&] { return all_at_least(h, h->arrive(), s); }, timeout_s)) {
    errno = ETIMEDOUT;
    return fail(-2, "shm collective: a peer never arrived");
  }
  for (int j = 0; j < h->world; ++j) {
    SlotHead pj;
    std::memcpy(&pj, h->slot(j), sizeof(pj));
    out_chunk[j] = pj.chunk_bytes;
    out_total[j] = pj.total_bytes;
    if (out != nullptr && pj.chunk_bytes) std::memcpy(out + (size_t)j * h->slot_bytes, h->slot(j) + sizeof(pj), pj.chunk_bytes);
  }
  h->depart()[h->rank].v.store(s, std::memory_order_release);
  return 0;
}

// Sizes first: every rank learns every rank's message length (so the caller can allocate).  Returns 0 / negative.
DETB200_HOST_API int detb200_shm_exchange_sizes(void* handle, uint64_t my_bytes, uint64_t* sizes_out, double timeout_s) {
  auto* h = static_cast<Handle*>(handle);
  uint64_t chunk[1024];
  if (h->world > 1024) return fail(-1, "world too large");
  return round_trip(h, nullptr, 0, my_bytes, nullptr, chunk, sizes_out, timeout_s);
}

// Allgather of variable-length byte strings.  `sizes` must come from detb200_shm_exchange_sizes of this same call; rank j's
// bytes land at out + offsets[j].  Messages longer than a slot stream through it in ceil(max / slot) rounds.
DETB200_HOST_API int detb200_shm_allgatherv(void* handle, const void* data, const uint64_t* sizes, const uint64_t* offsets, void* out,
                                            double timeout_s) {
  auto* h = static_cast<Handle*>(handle);
  if (h->world > 1024) return fail(-1, "world too large");
  uint64_t maxb = 0;
  for (int j = 0; j < h->world; ++j) maxb = sizes[j] > maxb ? sizes[j] : maxb;
  const uint64_t mine = sizes[h->rank];
  uint64_t chunk[1024], total[1024];
  uint8_t* stage = new uint8_t[(size_t)h->world * h->slot_bytes];
  int rc = 0;
  for (uint64_t off = 0; off < maxb && rc == 0; off += h->slot_bytes) {
    const uint64_t c = off < mine ? (mine - off < h->slot_bytes ? mine - off : h->slot_bytes) : 0;
    rc = round_trip(h, static_cast<const uint8_t*>(data) + off, c, mine, stage, chunk, total, timeout_s);
    if (rc) break;
    for (int j = 0; j < h->world; ++j) {
      if (total[j] != sizes[j]) {
        errno = 0;
        rc = fail(-3, "shm collective: ranks disagree about the call sequence");
        break;
      }
      if (chunk[j]) std::memcpy(static_cast<uint8_t*>(out) + offsets[j] + off, stage + (size_t)j * h->slot_bytes, chunk[j]);
    }
  }
  delete[] stage;
  return rc;
}

DETB200_HOST_API int detb200_shm_barrier(void* handle, double timeout_s) {
  auto* h = static_cast<Handle*>(handle);
  uint64_t chunk[1024], total[1024];
  if (h->world > 1024) return fail(-1, "world too large");
  return round_trip(h, nullptr, 0, 0, nullptr, chunk, total, timeout_s);
}

// Test hook: another rank's handle on the SAME mapping as `handle` (threads of one process).  ThreadSanitizer tracks
// virtual addresses, so only with a shared mapping can it see the protocol's data accesses and check that the
// arrive / depart release-acquire pairs really order them (csrc/host/shm_collective_test.cpp).
DETB200_HOST_API void* detb200_shm_clone_for_rank(void* handle, int rank) {
  auto* h0 = static_cast<Handle*>(handle);
  if (h0 == nullptr || rank < 0 || rank >= h0->world) return nullptr;
  auto* h = new Handle(*h0);
  h->rank = rank;
  h->round = 0;
  h->owner = false;
  return h;
}

DETB200_HOST_API void detb200_shm_free_clone(void* handle) { delete static_cast<Handle*>(handle); }
