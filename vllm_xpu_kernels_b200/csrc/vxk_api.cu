// Process-wide state of libvxk_b200.so and the ABI bookkeeping entry points.
#include "vxk_common.cuh"

namespace vxk {
thread_local char g_last_error[512] = {0};
std::atomic<uint64_t> g_launches{0};
// small-M GEMMs: programmatic dependent launch with the weight loads issued before griddepcontrol.wait
std::atomic<int> g_early_weight_loads{[] { const char* e = getenv("VXK_SKINNY_PDL"); return e ? atoi(e) : 1; }()};

int sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}
}  // namespace vxk

extern "C" int vxk_abi_version(void) { return VXK_ABI_VERSION; }
extern "C" const char* vxk_last_error(void) { return vxk::g_last_error; }
extern "C" uint64_t vxk_launch_count(void) { return vxk::g_launches.load(); }
extern "C" int vxk_set_early_weight_loads(int on) { return vxk::g_early_weight_loads.exchange(on ? 1 : 0); }
