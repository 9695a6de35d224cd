// Race check of the shared-memory collective under ThreadSanitizer:
//   g++ -std=c++17 -O1 -g -fsanitize=thread -pthread shm_collective.cpp shm_collective_test.cpp -lrt -o shm_tsan && ./shm_tsan
// N threads act as N ranks on ONE mapping (so TSan sees every access), run rounds of variable-length allgathers through a
// small slot (multi-round streaming) and verify the bytes.  Exit code 0 = data correct; TSan reports races on stderr and
// makes the exit code non-zero (halt_on_error is not needed: any report fails the run through TSAN_OPTIONS=exitcode).
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include <unistd.h>

extern "C" {
void* detb200_shm_open(const char* name, int rank, int world, size_t slot_bytes, double timeout_s);
void detb200_shm_close(void* h);
void* detb200_shm_clone_for_rank(void* h, int rank);
void detb200_shm_free_clone(void* h);
int detb200_shm_exchange_sizes(void* h, uint64_t my_bytes, uint64_t* sizes_out, double timeout_s);
int detb200_shm_allgatherv(void* h, const void* data, const uint64_t* sizes, const uint64_t* offsets, void* out, double timeout_s);
int detb200_shm_barrier(void* h, double timeout_s);
const char* detb200_shm_last_error(void);
}

static uint8_t byte_of(int rank, int round, uint64_t i) { return (uint8_t)(rank * 31 + round * 7 + i * 13); }

int main(int argc, char** argv) {
  const int world = argc > 1 ? std::atoi(argv[1]) : 4;
  const int rounds = argc > 2 ? std::atoi(argv[2]) : 300;
  const std::string name = "/detb200-tsan-" + std::to_string((long)getpid());
  void* h0 = detb200_shm_open(name.c_str(), 0, world, 256, 10.0);
  if (h0 == nullptr) {
    std::fprintf(stderr, "open failed: %s\n", detb200_shm_last_error());
    return 2;
  }
  std::vector<void*> hs(world);
  hs[0] = h0;
  for (int r = 1; r < world; ++r) hs[r] = detb200_shm_clone_for_rank(h0, r);
  std::vector<int> bad(world, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < world; ++r) {
    th.emplace_back([&, r] {
      std::vector<uint64_t> sizes(world), offs(world);
      for (int s = 0; s < rounds && !bad[r]; ++s) {
        const uint64_t mine = (uint64_t)((r * 97 + s * 53) % 1500);          // 0 .. 1499 bytes through 256-byte slots
        std::vector<uint8_t> data(mine);
        for (uint64_t i = 0; i < mine; ++i) data[i] = byte_of(r, s, i);
        if (detb200_shm_exchange_sizes(hs[r], mine, sizes.data(), 20.0)) { bad[r] = 1; break; }
        uint64_t total = 0;
        for (int j = 0; j < world; ++j) { offs[j] = total; total += sizes[j]; }
        std::vector<uint8_t> out(total ? total : 1);
        if (detb200_shm_allgatherv(hs[r], data.data(), sizes.data(), offs.data(), out.data(), 20.0)) { bad[r] = 2; break; }
        for (int j = 0; j < world && !bad[r]; ++j) {
          if (sizes[j] != (uint64_t)((j * 97 + s * 53) % 1500)) bad[r] = 3;
          for (uint64_t i = 0; i < sizes[j] && !bad[r]; ++i)
            if (out[offs[j] + i] != byte_of(j, s, i)) bad[r] = 4;
        }
        if ((s & 63) == 0 && detb200_shm_barrier(hs[r], 20.0)) bad[r] = 5;
      }
    });
  }
  for (auto& t : th) t.join();
  int rc = 0;
  for (int r = 0; r < world; ++r)
    if (bad[r]) {
      std::fprintf(stderr, "rank %d failed with code %d: %s\n", r, bad[r], detb200_shm_last_error());
      rc = 1;
    }
  for (int r = 1; r < world; ++r) detb200_shm_free_clone(hs[r]);
  detb200_shm_close(h0);
  if (rc == 0) std::printf("shm collective: %d ranks x %d rounds ok\n", world, rounds);
  return rc;
}
