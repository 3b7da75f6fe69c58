// Measured int8 tensor-pipe peak of this GPU (the denominator of every "tensor-bound" line): tcgen05.mma kind::i8,
// M = 128, N = 256, K = 32, operands in shared memory, accumulators in TMEM, issued back to back by one thread per CTA,
// one CTA per SM, for about a second.  Two TMEM accumulators alternate, a tcgen05.commit every 64 instructions bounds the
// number of instructions in flight.  Prints one JSON line: TOP/s (2 * M * N * K per instruction), cycles per instruction.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe_i8_peak tools/probe_i8_peak.cu
// usage: probe_i8_peak [seconds=1.0] [N=256]
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t a, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((a >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | ((uint64_t)1 << 46);
}

__global__ void __launch_bounds__(128, 1) peak(long long iters, int N, long long* cycles_out) {
    extern __shared__ __align__(1024) uint8_t smem[];      // A: 128 x 32 B, B: N x 32 B (K-major core matrices, contents irrelevant)
    __shared__ __align__(8) unsigned long long bar[2];     // alternate: a barrier never runs more than one phase ahead of its waiter
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (128 + 256) * 32; i += 128) smem[i] = (uint8_t)(i * 7 + 1);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[1])));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(s32(&tmem_base_s)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_base_s;
    if (tid == 0) {
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint64_t da = desc(s32(smem), 128, 256), db = desc(s32(smem) + 128 * 32, 128, 256);
        const long long t0 = clock64();
        auto wait = [&](long long k) {       // completion of the commit of batch k (barrier k & 1, its (k >> 1)-th phase)
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(s32(&bar[k & 1])), "r"((uint32_t)((k >> 1) & 1)) : "memory");
        };
        for (long long it = 0; it < iters; it++) {
#pragma unroll 8
            for (int j = 0; j < 64; j++)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                             ::"r"(tmem + (uint32_t)((j & 1) * 256)), "l"(da), "l"(db), "r"(idesc), "r"(1u), "r"(0u) : "memory");
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];" ::"r"(s32(&bar[it & 1])) : "memory");
            if (it >= 1) wait(it - 1);       // at most 128 instructions in flight
        }
        if (iters > 0) wait(iters - 1);
        cycles_out[blockIdx.x] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main(int argc, char** argv) {
    const double seconds = argc > 1 ? atof(argv[1]) : 1.0;
    const int N = argc > 2 ? atoi(argv[2]) : 256;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long* dcyc;
    cudaMalloc(&dcyc, sms * sizeof(long long));
    const int smem = (128 + 256) * 32;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    // calibrate, then the long run
    long long iters = 2000;
    float ms = 0;
    for (int pass = 0; pass < 2; pass++) {
        cudaEventRecord(e0);
        peak<<<sms, 128, smem>>>(iters, N, dcyc);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e)); return 1; }
        cudaEventElapsedTime(&ms, e0, e1);
        if (pass == 0) iters = (long long)(iters * seconds * 1e3 / ms) + 1;
    }
    long long* hc = (long long*)malloc(sms * sizeof(long long));
    cudaMemcpy(hc, dcyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double cyc = 0;
    for (int i = 0; i < sms; i++) cyc += (double)hc[i];
    cyc /= sms;
    const double n_mma = (double)iters * 64 * sms;
    const double tops = n_mma * 2.0 * 128 * N * 32 / (ms * 1e-3) / 1e12;
    printf("{\"what\": \"tcgen05.mma kind::i8 M128 N%d K32 back to back, one issuing thread per SM\", \"sms\": %d, \"seconds\": %.3f, "
           "\"int8_tops\": %.1f, \"cycles_per_mma\": %.2f, \"sm_mhz_effective\": %.0f}\n",
           N, sms, ms * 1e-3, tops, cyc / ((double)iters * 64), cyc / (ms * 1e-3) / 1e6);
    return 0;
}
