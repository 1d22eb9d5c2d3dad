// common.cuh -- error plumbing, launch accounting and small helpers shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <mutex>
#include <string>

#include "../../include/fp8approx_b200.h"

namespace b200 {

extern std::atomic<uint64_t> g_launches;
void set_error(const std::string& msg);
void set_path(const char* name);
int check_device();   // B200_OK, or B200_ERR_NO_DEVICE (there is no CPU fallback)

inline int cuda_fail(cudaError_t e, const char* what)
{
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    return B200_ERR_CUDA;
}

#define B200_CUDA_TRY(expr)                                          \
    do {                                                             \
        cudaError_t _e = (expr);                                     \
        if (_e != cudaSuccess) return ::b200::cuda_fail(_e, #expr);  \
    } while (0)

#define B200_LAUNCH_CHECK(name)                                          \
    do {                                                                 \
        ::b200::g_launches.fetch_add(1, std::memory_order_relaxed);      \
        cudaError_t _e = cudaGetLastError();                             \
        if (_e != cudaSuccess) return ::b200::cuda_fail(_e, name);       \
    } while (0)

// Per-device one-time setup (opt-in shared-memory sizes apply to the CURRENT device only, and the SM count belongs
// to it): `init` runs once per device, under a mutex, and the device is marked done only if it returned B200_OK.
struct PerDeviceOnce {
    static constexpr int kMaxDev = 64;
    std::mutex mu;
    bool done[kMaxDev] = {};
    int sms[kMaxDev] = {};
    template <class Init>
    int get(int* sms_out, Init init)
    {
        int dev = 0;
        B200_CUDA_TRY(cudaGetDevice(&dev));
        if (dev < 0 || dev >= kMaxDev) { set_error("device index out of range"); return B200_ERR_ARGUMENT; }
        std::lock_guard<std::mutex> lk(mu);
        if (!done[dev]) {
            int n = 0;
            B200_CUDA_TRY(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
            if (n <= 0) n = 148;
            if (int rc = init()) return rc;
            sms[dev] = n;
            done[dev] = true;
        }
        *sms_out = sms[dev];
        return B200_OK;
    }
};

// NVTX range around an entry point (tracing): header-only nvtx3, a no-op unless a tool (nsys, ncu
// --nvtx) is attached; the ranges name the C-ABI call a kernel belongs to in a timeline.
struct NvtxRange {
    explicit NvtxRange(const char* name);
    ~NvtxRange();
};

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

}  // namespace b200
