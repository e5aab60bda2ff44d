// Library-wide state of libelephas_b200.so: thread-local error string, launch counter, version.
#include "../include/elephas_b200.h"
#include "host.h"

namespace eb {

char* error_buffer() {
  static thread_local char buf[1024] = {0};
  return buf;
}

std::atomic<uint64_t>& launch_counter() {
  static std::atomic<uint64_t> c{0};
  return c;
}

}  // namespace eb

extern "C" const char* eb_last_error(void) { return eb::error_buffer(); }
extern "C" int eb_version(void) { return 100; }
extern "C" uint64_t eb_launch_count(void) { return eb::launch_counter().load(); }
