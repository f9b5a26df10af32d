// capi.cu -- library-wide state of libdcvc_b200 (error string, launch counter, version).
#include "common.cuh"

thread_local char g_dcvc_err[512] = "";
std::atomic<int64_t> g_dcvc_launches{0};
std::atomic<uint32_t> g_dcvc_once[DCVC_MAX_DEVICES];
int g_dcvc_num_sms[DCVC_MAX_DEVICES];
std::mutex g_dcvc_once_mu;

extern "C" const char *dcvc_last_error(void) { return g_dcvc_err; }
extern "C" int dcvc_version(void) { return 100; }
extern "C" int64_t dcvc_launch_count(void) { return g_dcvc_launches.load(); }
// kernels launched by replaying a CUDA graph that was recorded from this library's launches: the
// recorder (dcvc_b200/graphs.py) reports them, so that the counter keeps meaning "kernels run"
extern "C" void dcvc_launch_count_add(int64_t n) { g_dcvc_launches.fetch_add(n); }
