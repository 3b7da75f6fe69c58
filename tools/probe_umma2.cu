// Probe: the depthwise 64-tap temporal conv (layer 2) as tcgen05.mma with a HANKEL A operand described by
// overlapping descriptor strides -- no im2col copy:
//   A[m][k] = row[16 m + k]            K-major, no swizzle: core matrix = 8 rows x 16 B, rows 16 B apart (= the
//                                      byte array itself), next core matrix along K = +16 B (LBO = 16 B,
//                                      overlapping), next 8 rows = +128 B (SBO)
//   B[k][n] = w[k - n - 1]             Toeplitz of the 64 taps, [16][96] K-major core matrices
//   D[m][n] = sum_tap w[tap] row[16 m + n + 1 + tap]     (M = 128, N = 16, K = 3 x 32)
// and tcgen05.ld.32x32b.x16: thread i of warp w <- TMEM lane 32 w + i, registers = 16 consecutive columns.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe_umma2 tools/probe_umma2.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t a, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((a >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | ((uint64_t)1 << 46);
}

__global__ void __launch_bounds__(128, 1) probe(const int8_t* __restrict__ row, const int8_t* __restrict__ taps, int* __restrict__ out) {
    __shared__ __align__(128) uint8_t rs[2048 + 128];
    __shared__ __align__(128) uint8_t bs[16 * 96];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 2048 + 128; i += 128) rs[i] = (uint8_t)row[i];
    for (int i = tid; i < 16 * 96; i += 128) {
        const int n = i / 96, k = i % 96, tap = k - n - 1;
        bs[(n >> 3) * 768 + (k >> 4) * 128 + (n & 7) * 16 + (k & 15)] = (tap >= 0 && tap < 64) ? (uint8_t)taps[tap] : 0;
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(s32(&tmem_base_s)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_base_s;
    if (tid == 0) {
        // S32, int8 x int8, A K-major (bit 15 = 0), B K-major, N = 16, M = 128
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
        for (int ks = 0; ks < 3; ks++) {
            const uint64_t da = desc(s32(rs) + 32 * ks, 16, 128);
            const uint64_t db = desc(s32(bs) + 256 * ks, 128, 768);
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                         ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"((uint32_t)(ks > 0)), "r"(0u) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];" ::"r"(s32(&bar)) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(s32(&bar)));
    asm volatile("tcgen05.fence::after_thread_sync;");
    uint32_t r[16];
    const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(ta));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int n = 0; n < 16; n++) out[tid * 16 + n] = (int)r[n];
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem));
}

int main() {
    int8_t hr[2048 + 128], ht[64];
    srand(2);
    for (int i = 0; i < 2048 + 128; i++) hr[i] = (int8_t)(rand() % 255 - 127);
    for (int i = 0; i < 64; i++) ht[i] = (int8_t)(rand() % 255 - 127);
    int8_t *dr, *dt; int* dout;
    cudaMalloc(&dr, sizeof(hr)); cudaMalloc(&dt, 64); cudaMalloc(&dout, 128 * 16 * 4);
    cudaMemcpy(dr, hr, sizeof(hr), cudaMemcpyHostToDevice); cudaMemcpy(dt, ht, 64, cudaMemcpyHostToDevice);
    probe<<<1, 128>>>(dr, dt, dout);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel err=%d (%s)\n", (int)e, cudaGetErrorString(e));
    int ho[128 * 16];
    cudaMemcpy(ho, dout, sizeof(ho), cudaMemcpyDeviceToHost);
    int bad = 0, first = -1;
    for (int m = 0; m < 128; m++)
        for (int n = 0; n < 16; n++) {
            int ref = 0;
            for (int tap = 0; tap < 64; tap++) ref += (int)ht[tap] * (int)hr[16 * m + n + 1 + tap];
            if (ho[m * 16 + n] != ref) { bad++; if (first < 0) first = m * 16 + n; }
        }
    printf("mismatches=%d of 2048, first=%d\n", bad, first);
    if (bad) for (int n = 0; n < 4; n++) printf("  m=%d n=%d got=%d\n", first / 16, n, ho[(first / 16) * 16 + n]);
    return 0;
}
