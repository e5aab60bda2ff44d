// ThreadSanitizer harness for the host thread pool and the fp32 -> bf16 row conversion (csrc/host_stage.cu, host_cast.cpp):
// several caller threads convert batches of different shapes and thread counts concurrently (the pool grows under them) and
// every result is compared with a scalar conversion. Built and run by tests/utils/test_host_pool_tsan.py:
//   g++ -std=c++17 -O1 -g -fsanitize=thread -x c++ csrc/host_stage.cu csrc/host_cast.cpp tests/native/tsan_host_pool.cpp
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

extern "C" int eb_host_f32_to_bf16(const float* src, int64_t ld_src, void* dst, int64_t ld_dst, int64_t rows, int32_t cols,
                                   int32_t threads);

namespace eb {  // the two symbols host.h expects from the rest of the library
char* error_buffer() {
  static thread_local char buf[1024];
  return buf;
}
std::atomic<uint64_t>& launch_counter() {
  static std::atomic<uint64_t> c{0};
  return c;
}
}  // namespace eb

static uint16_t reference_cast(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7FFFFFFFu) > 0x7F800000u) return 0x7FFF;
  return static_cast<uint16_t>((u + 0x7FFFu + ((u >> 16) & 1u)) >> 16);
}

int main() {
  std::atomic<int> bad{0};
  auto caller = [&](int id) {
    std::mt19937 rng(1234 + id);
    std::normal_distribution<float> dist(0.f, 3.f);
    for (int it = 0; it < 40; ++it) {
      const int rows = 65 + static_cast<int>(rng() % 700), cols = 8 + static_cast<int>(rng() % 300);
      const int ld = (cols + 7) / 8 * 8, threads = 1 + static_cast<int>(rng() % 8);
      std::vector<float> src(static_cast<size_t>(rows) * cols);
      for (auto& v : src) v = dist(rng);
      std::vector<uint16_t> dst(static_cast<size_t>(rows) * ld + 32, 0xABCD);
      uint16_t* out = reinterpret_cast<uint16_t*>((reinterpret_cast<uintptr_t>(dst.data()) + 63) & ~uintptr_t(63));
      if (eb_host_f32_to_bf16(src.data(), cols, out, ld, rows, cols, threads) != 0) bad++;
      for (int r = 0; r < rows; ++r) {
        for (int c = 0; c < cols; ++c)
          if (out[static_cast<size_t>(r) * ld + c] != reference_cast(src[static_cast<size_t>(r) * cols + c])) bad++;
        for (int c = cols; c < ld; ++c)
          if (out[static_cast<size_t>(r) * ld + c] != 0) bad++;
      }
    }
  };
  std::vector<std::thread> callers;
  for (int i = 0; i < 4; ++i) callers.emplace_back(caller, i);
  for (auto& t : callers) t.join();
  printf("mismatches: %d\n", bad.load());
  return bad.load() ? 1 : 0;
}
