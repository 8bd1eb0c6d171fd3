#include "common.h"

namespace b200 {

char* error_buffer() {
    static thread_local char buf[512] = {0};
    return buf;
}

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

int sm_count_cached() {
    static thread_local int cached_dev = -1;
    static thread_local int cached_sms = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached_dev = dev;
        cached_sms = n;
    }
    return cached_sms;
}

}  // namespace b200

extern "C" const char* b200_last_error(void) { return b200::error_buffer(); }
extern "C" int b200_version(void) { return 100; }
extern "C" int b200_sm_count(void) { return b200::sm_count_cached(); }

// Event trace of the decode kernels (csrc/trace.cuh): buf = device memory of 8 * (1 + 2 * cap) bytes, zeroed by the
// caller; NULL switches the trace off again.
extern "C" int b200_trace_set_gemm(void*, int);
extern "C" int b200_trace_set_attn(void*, int);
extern "C" int b200_trace_set(void* buf, int32_t cap) {
    if (b200_trace_set_gemm(buf, cap) != 0 || b200_trace_set_attn(buf, cap) != 0)
        return b200::fail(B200_ECUDA, "b200_trace_set: cudaMemcpyToSymbol failed");
    return B200_OK;
}
