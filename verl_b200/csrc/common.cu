#include "common.h"

#include <atomic>
#include <mutex>

namespace vpg {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
    static std::mutex mu;
    static int cache[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    std::lock_guard<std::mutex> lk(mu);
    if (cache[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev] = n;
    }
    return cache[dev];
}

}  // namespace vpg

extern "C" {

const char* vpg_version(void) { return "vpg_b200 0.1.0 (sm_100a)"; }
const char* vpg_last_error(void) { return vpg::g_err; }
int64_t vpg_launch_count(void) { return vpg::g_launches.load(std::memory_order_relaxed); }
int vpg_stream_synchronize(void* stream) {
    VPG_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    return VPG_OK;
}

}  // extern "C"
