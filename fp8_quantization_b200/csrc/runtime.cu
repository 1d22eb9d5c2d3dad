// runtime.cu -- library state (last error, launch counter), device check, and the table selector
// A5 get_error_table_NN (approx/approx_matmul_whole_v9.py:555-592).
#include <map>
#include <mutex>
#include <string.h>

#include "common.cuh"
#include "comp_tables.inc"
#include "comp_tables_v5.inc"
#include "tables.h"

#include <cstddef>
static_assert(sizeof(b200_approx_params) == 56 && offsetof(b200_approx_params, ofuf_mode) == 24 &&
                  offsetof(b200_approx_params, prepared_b_ready) == 26 && offsetof(b200_approx_params, custom_table) == 32 &&
                  offsetof(b200_approx_params, prepared_b) == 40 && offsetof(b200_approx_params, prepared_b_bytes) == 48,
              "b200_approx_params layout is part of the C ABI (fp8_quantization_b200/_lib.py mirrors it)");

#include <nvtx3/nvToolsExt.h>

namespace b200 {

NvtxRange::NvtxRange(const char* name) { nvtxRangePushA(name); }
NvtxRange::~NvtxRange() { nvtxRangePop(); }

std::atomic<uint64_t> g_launches{0};
static std::mutex g_err_mu;
static std::string g_err = "";
static thread_local char g_err_buf[512];
static const char* g_path = "none";

void set_error(const std::string& msg)
{
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_err = msg;
}
void set_path(const char* name) { g_path = name; }

int check_device()
{
    static int cached = 1;   // 1 = unknown
    if (cached != 1) return cached;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error("no CUDA device visible: libfp8approx_b200 has no CPU fallback");
        (void)cudaGetLastError();
        return B200_ERR_NO_DEVICE;   // not cached: a later call may see a device
    }
    int dev = 0, major = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (major != 10) {
        set_error("device is not sm_100: this library ships sm_100a code only");
        return B200_ERR_NO_DEVICE;
    }
    cached = B200_OK;
    return cached;
}

// Table choice of get_error_table_NN (v9 :555-592).  Returns a pointer to 4^M int8 entries, or
// nullptr for the all-zero table.
int select_table(int E, int M, int withComp, int dnsmp, int allow_unpinned, const signed char** out)
{
    *out = nullptr;
    if (E == 4 && M == 3) {
        if (!withComp) *out = kCompTable0;
        return B200_OK;
    }
    if (E == 3 && M == 4) {
        if (withComp) {
            if (dnsmp == 3) *out = kCompTable1;
            else if (dnsmp >= 4) *out = nullptr;
            else { set_error("UnboundLocalError: error_table_NN (E3M4, withComp, dnsmp_factor < 3)"); return B200_ERR_UNBOUND_TABLE; }
        } else {
            *out = kCompTable2;
        }
        return B200_OK;
    }
    if (E == 2 && M == 5) {
        if (withComp) {
            if (dnsmp == 3) *out = kCompTable3;
            else if (dnsmp == 4) *out = kCompTable4;
            else if (dnsmp == 5) *out = nullptr;
            else { set_error("UnboundLocalError: error_table_NN (E2M5, withComp, dnsmp_factor not in 3..5)"); return B200_ERR_UNBOUND_TABLE; }
        } else {
            *out = kCompTable5;
        }
        return B200_OK;
    }
    if (allow_unpinned && E >= 1 && M >= 1 && M <= 5 && E + M <= 7) return B200_OK;   // zero table, no reference answer
    set_error("Invalid combination of expo_width and mant_width");
    return B200_ERR_INVALID_FORMAT;
}

// get_comp_table_NN of the v5 integer-add multiplier (approx/approx_matmul_whole_v5.py:516-542); nullptr = zeros
// (without withComp v5 assigns a zero table and then raises: the opt-in mode uses the zero table).
int select_table_v5(int E, int M, int withComp, int dnsmp, const signed char** out)
{
    *out = nullptr;
    const bool known = (E == 4 && M == 3) || (E == 3 && M == 4) || (E == 2 && M == 5);
    if (!known) { set_error("Invalid combination of expo_width and mant_width"); return B200_ERR_INVALID_FORMAT; }
    if (!withComp) return B200_OK;
    if (E == 4) { *out = kCompV5Table0; return B200_OK; }
    if (E == 3) {
        if (dnsmp == 3) *out = kCompV5Table1;
        else if (dnsmp >= 4) *out = kCompV5Table2;
        else { set_error("UnboundLocalError: comp_table_NN (E3M4, withComp, dnsmp_factor < 3)"); return B200_ERR_UNBOUND_TABLE; }
        return B200_OK;
    }
    if (dnsmp == 3) *out = kCompV5Table3;
    else if (dnsmp == 4) *out = kCompV5Table4;
    else if (dnsmp >= 5) *out = kCompV5Table5;
    else { set_error("UnboundLocalError: comp_table_NN (E2M5, withComp, dnsmp_factor < 3)"); return B200_ERR_UNBOUND_TABLE; }
    return B200_OK;
}

const signed char* intern_table(const signed char* host, int M)
{
    static std::mutex mu;
    static std::map<std::string, std::string*> pool;
    const size_t n = (size_t)(1 << M) * (1 << M);
    std::string key(reinterpret_cast<const char*>(host), n);
    std::lock_guard<std::mutex> lk(mu);
    auto it = pool.find(key);
    if (it == pool.end()) it = pool.emplace(key, new std::string(key)).first;
    return reinterpret_cast<const signed char*>(it->second->data());
}

}  // namespace b200

using namespace b200;

extern "C" const char* b200_version(void) { return "fp8approx_b200 0.1 (sm_100a)"; }

extern "C" const char* b200_last_error(void)
{
    std::lock_guard<std::mutex> lk(g_err_mu);
    strncpy(g_err_buf, g_err.c_str(), sizeof(g_err_buf) - 1);
    g_err_buf[sizeof(g_err_buf) - 1] = 0;
    return g_err_buf;
}

extern "C" uint64_t b200_kernel_launches(void) { return g_launches.load(); }
extern "C" const char* b200_last_path(void) { return g_path; }

extern "C" int b200_comp_table_v5(int E, int M, int withComp, int dnsmp, int8_t* host_out)
{
    const signed char* t = nullptr;
    int rc = select_table_v5(E, M, withComp, dnsmp, &t);
    if (rc != B200_OK) return rc;
    if (!host_out) { set_error("b200_comp_table_v5: host_out is NULL"); return B200_ERR_ARGUMENT; }
    const int n = (1 << M) * (1 << M);
    if (t) memcpy(host_out, t, n);
    else memset(host_out, 0, n);
    return B200_OK;
}

extern "C" int b200_error_table(int E, int M, int withComp, int dnsmp, int allow_unpinned, int8_t* host_out)
{
    const signed char* t = nullptr;
    int rc = select_table(E, M, withComp, dnsmp, allow_unpinned, &t);
    if (rc != B200_OK) return rc;
    const int n = (1 << M) * (1 << M);
    if (!host_out) { set_error("b200_error_table: host_out is NULL"); return B200_ERR_ARGUMENT; }
    if (t) memcpy(host_out, t, n);
    else memset(host_out, 0, n);
    return B200_OK;
}
