// common.cu — process-wide state of libb200seq.so (error string, launch counter, version).
#include "common.h"
#include "../include/b200seq.h"

namespace b200 {
thread_local char g_err[512] = {0};
std::atomic<long long> g_launches{0};
}  // namespace b200

extern "C" const char* b200_last_error(void) { return b200::g_err; }
extern "C" int b200_version(void) { return 100; }
extern "C" long long b200_launch_count(void) { return b200::g_launches.load(); }
extern "C" void b200_reset_launch_count(void) { b200::g_launches.store(0); }
// replays of a captured CUDA graph re-issue the launches that were counted while capturing
extern "C" void b200_add_launch_count(long long n) { b200::g_launches.fetch_add(n); }
