// Library-level entry points: version, status strings, launch accounting.
#include <atomic>
#include <cstdlib>

#include "common.cuh"

namespace hymeta {
thread_local int g_last_cuda_error = 0;
static std::atomic<int64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
bool pdl_enabled() {
  static const bool on = [] {
    const char* s = getenv("HYMETA_PDL");
    return !(s && s[0] == '0');
  }();
  return on;
}
}  // namespace hymeta

extern "C" int hymeta_abi_version(void) { return HYMETA_ABI_VERSION; }

extern "C" const char* hymeta_status_string(int status) {
  switch (status) {
    case HYMETA_OK: return "ok";
    case HYMETA_ERR_INVALID_ARG: return "invalid argument (null pointer, bad size or stride)";
    case HYMETA_ERR_UNSUPPORTED: return "unsupported geometry (head_dim must be 128; see hymeta_b200.h)";
    case HYMETA_ERR_WORKSPACE: return "workspace too small";
    case HYMETA_ERR_CUDA: return "CUDA launch failed";
    case HYMETA_ERR_NO_DEVICE: return "no sm_100 device or driver entry point unavailable";
    case HYMETA_ERR_NO_SLOT: return "no free state slot: more live sequences than max_num_seqs";
    default: return "unknown status";
  }
}

extern "C" int hymeta_last_cuda_error(void) { return hymeta::g_last_cuda_error; }
extern "C" int64_t hymeta_launch_count(void) { return hymeta::g_launches.load(); }
