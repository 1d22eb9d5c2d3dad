// probe.cu -- measures the CUDA-core integer issue peak (the roofline denominator of the approx GEMM,
// "INT_PEAK is not in MEASURED_PEAKS.json: measure once with a register-only
// IADD3/LOP3 saturation kernel, all SMs, >= 4 independent chains per thread").
#include "common.cuh"

namespace b200 {

// 8 independent chains per thread, alternating IADD3 and LOP3 (both alu-pipe instructions).
__global__ void __launch_bounds__(256) int_peak_kernel(int iters, uint32_t seed, uint32_t* __restrict__ sink)
{
    uint32_t x[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) x[c] = seed + threadIdx.x * 8 + c;
    const uint32_t k1 = seed | 1u, k2 = seed ^ 0x9E3779B9u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                asm volatile("add.u32 %0, %0, %1;" : "+r"(x[c]) : "r"(k1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[c]) : "r"(k2), "r"(k1));
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) r ^= x[c];
    if (r == 0x12345678u) sink[0] = r;   // practically never; keeps the chains alive
}

// Shared-memory lookup peak: the bound of the table GEMM (one 16-bit lookup per MAC).  Every lane reads its own
// bank (byte address 4 * lane + a multiple of 128), so each warp-wide LD.SHARED.U16 is ONE conflict-free wavefront;
// 16 independent loads per iteration per thread keep the LSU pipe full, the XOR per load keeps them alive.
__global__ void __launch_bounds__(256) lds_peak_kernel(int iters, uint32_t seed, uint32_t* __restrict__ sink)
{
    __shared__ __align__(16) uint16_t tab[16 * 1024];   // 32 KB
    for (int i = threadIdx.x; i < 16 * 1024; i += 256) tab[i] = (uint16_t)(i * 2654435761u + seed);
    __syncthreads();
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(tab) + 4u * (threadIdx.x & 31) + 2048u * (threadIdx.x >> 5);
    uint32_t acc[4] = {seed, seed + 1, seed + 2, seed + 3};
    for (int it = 0; it < iters; ++it) {
#define B200_LDS_PROBE(U)                                                                   \
        {                                                                                   \
            uint32_t v;                                                                     \
            asm volatile("ld.volatile.shared.u16 %0, [%1+%2];" : "=r"(v) : "r"(base), "n"(128 * (U))); \
            acc[(U) & 3] ^= v;                                                              \
        }
        B200_LDS_PROBE(0) B200_LDS_PROBE(1) B200_LDS_PROBE(2) B200_LDS_PROBE(3) B200_LDS_PROBE(4) B200_LDS_PROBE(5)
        B200_LDS_PROBE(6) B200_LDS_PROBE(7) B200_LDS_PROBE(8) B200_LDS_PROBE(9) B200_LDS_PROBE(10) B200_LDS_PROBE(11)
        B200_LDS_PROBE(12) B200_LDS_PROBE(13) B200_LDS_PROBE(14) B200_LDS_PROBE(15)
#undef B200_LDS_PROBE
    }
    const uint32_t r = acc[0] ^ acc[1] ^ acc[2] ^ acc[3];
    if (r == 0x12345678u) sink[0] = r;
}

}  // namespace b200

using namespace b200;

// Conflict-free 16-bit shared-memory lookups per second over all SMs (T lookups/s).  Synchronous; a measurement
// tool (bench.py's roofline denominator for the table GEMM), not part of the hot path.
extern "C" int b200_lds_peak_probe(int iters, double* tlookups_out, double* ms_out)
{
    if (int rc = check_device()) return rc;
    if (iters < 1 || !tlookups_out) { set_error("b200_lds_peak_probe: bad argument"); return B200_ERR_ARGUMENT; }
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t* sink = nullptr;
    B200_CUDA_TRY(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 4;   // 4 x 8 warps resident per SM
    lds_peak_kernel<<<blocks, 256>>>(iters / 4 + 1, 7u, sink);   // warm-up
    double best_ms = 1e30;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        lds_peak_kernel<<<blocks, 256>>>(iters, 7u + rep, sink);
        cudaEventRecord(e1);
        B200_CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best_ms) best_ms = ms;
    }
    g_launches.fetch_add(4);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    *tlookups_out = (double)blocks * 256.0 * (double)iters * 16.0 / (best_ms * 1e-3) / 1e12;
    if (ms_out) *ms_out = best_ms;
    return B200_OK;
}

// Runs the saturation kernel on every SM and reports integer lane-ops per second (Tiop/s).
// Synchronous (it is a measurement tool, not part of the hot path).
extern "C" int b200_int_peak_probe(int iters, double* tiops_out, double* ms_out)
{
    if (int rc = check_device()) return rc;
    if (iters < 1 || !tiops_out) { set_error("b200_int_peak_probe: bad argument"); return B200_ERR_ARGUMENT; }
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t* sink = nullptr;
    B200_CUDA_TRY(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8;
    int_peak_kernel<<<blocks, 256>>>(iters / 4 + 1, 12345u, sink);   // warm-up
    double best_ms = 1e30;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        int_peak_kernel<<<blocks, 256>>>(iters, 12345u + rep, sink);
        cudaEventRecord(e1);
        B200_CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best_ms) best_ms = ms;
    }
    g_launches.fetch_add(4);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    const double ops = (double)blocks * 256.0 * (double)iters * 8.0 * 8.0 * 2.0;
    *tiops_out = ops / (best_ms * 1e-3) / 1e12;
    if (ms_out) *ms_out = best_ms;
    return B200_OK;
}
