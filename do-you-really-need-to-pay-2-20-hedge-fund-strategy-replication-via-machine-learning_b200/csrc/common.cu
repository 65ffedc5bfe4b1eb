// Error plumbing, device queries and the window index tables of libhfrep.
#include "common.cuh"

#include <atomic>

namespace hfrep {

char* err_buf() {
    static thread_local char buf[512] = "";
    return buf;
}

int set_err(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}

int sm_count() {
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cached[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

}  // namespace hfrep

extern "C" const char* hfrep_last_error(void) { return hfrep::err_buf(); }
extern "C" int hfrep_version(void) { return 200; }

// Window bookkeeping of the reference, bit-exact.
extern "C" int hfrep_window_index(int kind, int T, int window, int32_t* starts, int32_t* lens, int cap) {
    HFREP_REQUIRE(T >= 0 && window >= 0 && cap >= 0, "window_index: negative argument");
    int n = 0;
    switch (kind) {
        case 0:   // rolling OLS windows: for i in range(len(x_test) - window)   Autoencoder_encapsulate.py:148-156
            for (int i = 0; i < T - window; ++i, ++n)
                if (n < cap) { if (starts) starts[n] = i; if (lens) lens[n] = window; }
            break;
        case 1:   // expanding prefixes: for i in range(2, len(x_test)): x_test[:i]   Autoencoder_encapsulate.py:117,126
            for (int i = 2; i < T; ++i, ++n)
                if (n < cap) { if (starts) starts[n] = 0; if (lens) lens[n] = i; }
            break;
        case 2:   // ex-post covariance windows: for i in range(1, len(factor_etf) - window)   helper.py:120-121
            for (int i = 1; i < T - window; ++i, ++n)
                if (n < cap) { if (starts) starts[n] = i; if (lens) lens[n] = window; }
            break;
        default:
            return hfrep::set_err(HFREP_ERR_ARG, "window_index: kind=%d (0 rolling, 1 expanding, 2 expost)", kind);
    }
    return n;
}
