// Error state, device check and version of libdexgg.
#include "common.cuh"

#include <atomic>
#include <mutex>

namespace dexgg {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};
void count_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
unsigned long long launches() { return g_launches.load(std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static int check_device(int dev, int* sms) {
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, dev);
    if (e != cudaSuccess) {
        set_error("cudaGetDeviceProperties failed: %s", cudaGetErrorString(e));
        return DEXGG_E_CUDA;
    }
    if (sms) *sms = prop.multiProcessorCount;
    if (prop.major != 10) {
        set_error("device %d (%s) is sm_%d%d; libdexgg runs on sm_100 (B200) only and has no fallback", dev, prop.name,
                  prop.major, prop.minor);
        return DEXGG_E_ARCH;
    }
    return DEXGG_OK;
}

struct DevInfo {
    int rc = 1;  // 1 = unknown
    int sms = 0;
};
static DevInfo g_dev[64];
static std::mutex g_mu;

static DevInfo& current_info() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        static DevInfo bad;
        bad.rc = DEXGG_E_CUDA;
        set_error("no usable CUDA device (cudaGetDevice failed); libdexgg has no CPU fallback");
        return bad;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_dev[dev].rc == 1) g_dev[dev].rc = check_device(dev, &g_dev[dev].sms);
    return g_dev[dev];
}

int device_check_cached() {
    DevInfo& d = current_info();
    if (d.rc == DEXGG_E_ARCH) set_error("current device is not sm_100 (B200); libdexgg has no fallback");
    return d.rc;
}

int sm_count() {
    DevInfo& d = current_info();
    return d.sms > 0 ? d.sms : 148;
}

}  // namespace dexgg

extern "C" int dexgg_version(void) { return DEXGG_VERSION; }
extern "C" const char* dexgg_last_error(void) { return dexgg::g_err; }
extern "C" int dexgg_device_check(void) { return dexgg::device_check_cached(); }
extern "C" unsigned long long dexgg_launch_count(void) { return dexgg::launches(); }
