// Probe: layer 2 (depthwise 64-tap conv) with N = 32 outputs per tcgen05.mma row.
//   A[m][kk] = S[32 m + kk]     Hankel view of the byte array S with a ROW PITCH OF 32 BYTES: K-major SWIZZLE_32B descriptor
//                               (8 rows x 32 B atom, SBO = 256 B); k-step ks = start address + 32 ks (i.e. "the next row").
//                               The bytes are stored where an ABSOLUTE-address XOR (16-byte chunk bit 4 ^= address bit 7) puts them.
//   B'[n'][kk] = w[kk + n' - 32] Toeplitz with REVERSED output order (column n' = output 31 - n'), so that
//                               B'[n' + 16][kk - 16] = B'[n'][kk]: the core matrices overlap (LBO = 256 B, SBO = 128 B) and the
//                               operand takes 1792 bytes instead of 3072.
//   D[m][n'] = sum_tap w[tap] S[32 m + (31 - n') + 1 + tap]          (M = 128, N = 32, K = 3 x 32)
// Questions: (1) does the swizzle use absolute address bits (start addresses that are not 256-byte aligned, start + 32 ks)?
//            (2) do overlapping B core matrices work?  (3) tcgen05.ld.32x32b.x32.
// usage: probe_umma3   -> one line per (base offset, B overlap) combination
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe_umma3 tools/probe_umma3.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t a, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((a >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)layout << 61);
}

constexpr int kRows = 128, kS = 32 * kRows + 96;

__global__ void __launch_bounds__(128, 1) probe(const int8_t* __restrict__ S, const int8_t* __restrict__ taps, int* __restrict__ out,
                                                int base_off, int b_overlap, int use_base_offset_field) {
    __shared__ __align__(1024) uint8_t as[kS + 1024];
    __shared__ __align__(128) uint8_t bs[32 * 96];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t abase = s32(as) + base_off;
    for (int i = tid; i < kS; i += 128) {
        const uint32_t a = abase + i, p = a ^ (((a >> 7) & 1u) << 4);
        as[p - s32(as)] = (uint8_t)S[i];
    }
    for (int i = tid; i < 32 * 96; i += 128) bs[i] = 0;
    __syncthreads();
    for (int i = tid; i < 32 * 96; i += 128) {
        const int n = i / 96, k = i % 96, tap = k + n - 32;
        const uint8_t v = (tap >= 0 && tap < 64) ? (uint8_t)taps[tap] : 0;
        if (b_overlap) bs[(n >> 3) * 128 + (k >> 4) * 256 + (n & 7) * 16 + (k & 15)] = v;     // overlapping writes carry equal values
        else bs[(n >> 3) * 768 + (k >> 4) * 128 + (n & 7) * 16 + (k & 15)] = v;
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
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
        for (int ks = 0; ks < 3; ks++) {
            const uint32_t sa = abase + 32 * ks;
            uint64_t da = desc(sa, 16, 256, 6);                    // SWIZZLE_32B
            if (use_base_offset_field) da |= (uint64_t)((sa >> 7) & 7) << 49;
            const uint64_t db = b_overlap ? desc(s32(bs) + 512 * ks, 256, 128, 0) : desc(s32(bs) + 256 * ks, 128, 768, 0);
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
    uint32_t r[32];
    const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(ta));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int n = 0; n < 32; n++) out[tid * 32 + n] = (int)r[n];
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem));
}

int main() {
    static int8_t hs[kS], ht[64];
    srand(3);
    for (int i = 0; i < kS; i++) hs[i] = (int8_t)(rand() % 255 - 127);
    for (int i = 0; i < 64; i++) ht[i] = (int8_t)(rand() % 255 - 127);
    int8_t *ds, *dt; int* dout;
    cudaMalloc(&ds, kS); cudaMalloc(&dt, 64); cudaMalloc(&dout, 128 * 32 * 4);
    cudaMemcpy(ds, hs, kS, cudaMemcpyHostToDevice); cudaMemcpy(dt, ht, 64, cudaMemcpyHostToDevice);
    static int ho[128 * 32];
    const int offs[] = {0, 128, 256, 16, 48, 32, 64, 80, 144, 400, 912};
    for (int field = 0; field < 2; field++)
        for (int bo = 0; bo < 2; bo++)
            for (int oi = 0; oi < (int)(sizeof(offs) / sizeof(offs[0])); oi++) {
                cudaMemset(dout, 0xff, sizeof(ho));
                probe<<<1, 128>>>(ds, dt, dout, offs[oi], bo, field);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("base_off=%d b_overlap=%d field=%d kernel err=%d (%s)\n", offs[oi], bo, field, (int)e, cudaGetErrorString(e)); return 1; }
                cudaMemcpy(ho, dout, sizeof(ho), cudaMemcpyDeviceToHost);
                int bad = 0, first = -1;
                for (int m = 0; m < 128; m++)
                    for (int n = 0; n < 32; n++) {
                        int ref = 0;
                        for (int tap = 0; tap < 64; tap++) ref += (int)ht[tap] * (int)hs[32 * m + (31 - n) + 1 + tap];
                        if (ho[m * 32 + n] != ref) { bad++; if (first < 0) first = m * 32 + n; }
                    }
                printf("base_off=%4d b_overlap=%d base_offset_field=%d mismatches=%d of 4096 first=(m %d, n %d)\n", offs[oi], bo, field, bad,
                       first < 0 ? -1 : first / 32, first < 0 ? -1 : first % 32);
            }
    return 0;
}
