// Probe: layer 1 as ONE tcgen05.mma (kind::i8) tile on sm_100a.
//   D[128 samples x 16 filters] (int32, TMEM) = Xt[128 x 64 ch] * W1t[64 x 16]
//   A = the [64 ch][128 samples] TMA tile with the 128-byte swizzle, read as an MN-major operand
//       (canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units: K row = 128 contiguous bytes,
//        8 K rows = one 1024-byte swizzle atom, SBO = 1024 B)
//   B = W1 [16][64] K-major, no swizzle: 8x16-byte core matrices, LBO = 128 B (next 16 channels),
//       SBO = 512 B (next 8 filters)
// and the accumulator read back with tcgen05.ld.16x256b (expected: the m16n8 C-fragment layout of
// mma.sync: thread (g, q) gets rows g / g+8, columns 2q, 2q+1 of every 8-column block).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe_umma tools/probe_umma.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1)
probe(const __grid_constant__ CUtensorMap tm, const int8_t* __restrict__ w1, int* __restrict__ out, int idesc_override) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* xs = smem;                 // [64][128] swizzled tile, 8 KB
    uint8_t* ws = smem + 8192;          // W1 core matrices, 1 KB
    __shared__ __align__(8) unsigned long long bar_x, bar_mma;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar_x)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar_mma)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    // W1 [16][64] -> core-matrix layout
    for (int i = tid; i < 16 * 64; i += 128) {
        const int n = i >> 6, c = i & 63;
        ws[(n >> 3) * 512 + (c >> 4) * 128 + (n & 7) * 16 + (c & 15)] = (uint8_t)w1[i];
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(s32(&tmem_base_s)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic writes of ws -> async proxy (UMMA reads)
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_base_s;
    if (tid == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar_x)), "r"(8192));
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(s32(xs)), "l"(&tm), "r"(0), "r"(0), "r"(s32(&bar_x)) : "memory");
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(s32(&bar_x)));
        asm volatile("tcgen05.fence::after_thread_sync;");
        // instruction descriptor: S32 accumulate, A = B = signed int8, A MN-major, B K-major, N = 16, M = 128
        uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (0u << 16) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
        if (idesc_override) idesc = (uint32_t)idesc_override;
        for (int k = 0; k < 2; k++) {
            // A: start = tile + k * 32 rows * 128 B, SBO = 1024 B, LBO unused (one 128-byte block along M)
            const uint64_t da = (uint64_t)(((s32(xs) + k * 4096) >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
                                ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
            // B: start = ws + k * 2 core matrices along K, LBO = 128 B, SBO = 512 B, no swizzle
            const uint64_t db = (uint64_t)(((s32(ws) + k * 256) >> 4) & 0x3fff) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) |
                                ((uint64_t)1 << 46);
            const uint32_t acc = k > 0;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                         ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(0u) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];" ::"r"(s32(&bar_mma)) : "memory");
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(s32(&bar_mma)));
    }
    asm volatile("tcgen05.fence::after_thread_sync;");
    // warp w owns TMEM lanes 32 w .. 32 w + 31 = samples; two 16-lane halves, 16 columns = .x2
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        uint32_t r[8];
        const uint32_t taddr = tmem + ((uint32_t)(32 * warp + 16 * mt) << 16);
        asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        // assumed: block nt (8 columns) = regs 4 nt .. 4 nt + 3 = (row g, col 2q), (g, 2q+1), (g+8, 2q), (g+8, 2q+1)
#pragma unroll
        for (int nt = 0; nt < 2; nt++)
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int sample = 32 * warp + 16 * mt + g + 8 * (c >> 1), filt = 8 * nt + 2 * q + (c & 1);
                out[sample * 16 + filt] = (int)r[4 * nt + c];
            }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem));
}

// throughput: `reps` back-to-back tcgen05.mma (M = 128, N = n, K = 32, i8) from one thread, then one commit
__global__ void __launch_bounds__(128, 1) timing(int reps, int n, long long* cyc, int astep, int mode, int issuers) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 65536 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u * (i & 3);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
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
    if (issuers > 1) {
        __shared__ long long tt[8];
        if ((tid & 31) == 0 && warp < issuers) {
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
            const uint64_t da = (uint64_t)(((s32(smem) + 8192 * warp) >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
            const uint64_t db = (uint64_t)(((s32(smem) + 49152) >> 4) & 0x3fff) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | ((uint64_t)1 << 46);
            const uint32_t d = tmem + 64 * warp;
            const long long t0 = clock64();
            for (int i = 0; i < reps; i++)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                             ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(1u), "r"(0u) : "memory");
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];" ::"r"(s32(&bar)) : "memory");
            tt[warp] = clock64() - t0;
        }
        __syncthreads();
        if (tid == 0) {
            uint32_t done = 0;   // barrier was initialised with count 1: only valid to wait for the first commit; use the slowest issue time instead
            long long mx = 0;
            for (int i = 0; i < issuers; i++) mx = tt[i] > mx ? tt[i] : mx;
            cyc[0] = mx; cyc[1] = mx;
            (void)done;
        }
    } else
    if (tid == 0) {
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
        // mode 0: A = MN-major swizzled tile (layer 1); mode 1: A = K-major Hankel view, LBO 16 B / SBO 128 B (layer 2)
        const uint32_t idesc2 = mode ? (idesc & ~(1u << 15)) : idesc;
        const uint64_t da0 = mode ? ((uint64_t)((s32(smem) >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(128 >> 4) << 32) | ((uint64_t)1 << 46))
                                  : ((uint64_t)((s32(smem) >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61));
        const uint64_t db = (uint64_t)(((s32(smem) + 49152) >> 4) & 0x3fff) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | ((uint64_t)1 << 46);
        const long long t0 = clock64();
        for (int i = 0; i < reps; i++) {
            const uint32_t d = tmem + (uint32_t)((i & 3) * 64);          // 4 independent accumulators
            const uint64_t da = da0 + (uint64_t)((i & 7) * astep);       // A operand walks over 8 x astep x 16 bytes
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                         ::"r"(d), "l"(da), "l"(db), "r"(idesc2), "r"(1u), "r"(0u) : "memory");
        }
        const long long t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];" ::"r"(s32(&bar)) : "memory");
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(s32(&bar)));
        const long long t2 = clock64();
        cyc[0] = t1 - t0; cyc[1] = t2 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main(int argc, char** argv) {
    const int C = 64, L = 128;
    int8_t hx[C * L], hw[16 * C];
    srand(1);
    for (int i = 0; i < C * L; i++) hx[i] = (int8_t)(rand() % 255 - 127);
    for (int i = 0; i < 16 * C; i++) hw[i] = (int8_t)(rand() % 255 - 127);
    int8_t *dx, *dw; int* dout;
    cudaMalloc(&dx, sizeof(hx)); cudaMalloc(&dw, sizeof(hw)); cudaMalloc(&dout, 128 * 16 * 4);
    cudaMemcpy(dx, hx, sizeof(hx), cudaMemcpyHostToDevice); cudaMemcpy(dw, hw, sizeof(hw), cudaMemcpyHostToDevice);
    cudaMemset(dout, 0xff, 128 * 16 * 4);
    typedef CUresult (*enc_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
    CUtensorMap tm;
    cuuint64_t dims[2] = {L, C}; cuuint64_t strides[1] = {L}; cuuint32_t box[2] = {128, 64}; cuuint32_t estr[2] = {1, 1};
    CUresult r = ((enc_t)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, dx, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode=%d\n", (int)r);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
    probe<<<1, 128, 16384>>>(tm, dw, dout, argc > 1 ? (int)strtol(argv[1], 0, 0) : 0);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel err=%d (%s)\n", (int)e, cudaGetErrorString(e));
    int ho[128 * 16];
    cudaMemcpy(ho, dout, sizeof(ho), cudaMemcpyDeviceToHost);
    int bad = 0, first = -1;
    for (int t = 0; t < 128; t++)
        for (int f = 0; f < 16; f++) {
            int ref = 0;
            for (int c = 0; c < C; c++) ref += (int)hx[c * L + t] * (int)hw[f * C + c];
            if (ho[t * 16 + f] != ref) { bad++; if (first < 0) first = t * 16 + f; }
        }
    printf("mismatches=%d of 2048, first=%d\n", bad, first);
    if (bad) for (int i = 0; i < 8; i++) {
        int t = (first / 16), f = i; int ref = 0;
        for (int c = 0; c < C; c++) ref += (int)hx[c * L + t] * (int)hw[f * C + c];
        printf("  t=%d f=%d got=%d ref=%d\n", t, f, ho[t * 16 + f], ref);
    }
    {
        long long* dc; cudaMalloc(&dc, 16);
        cudaFuncSetAttribute(timing, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
        const int ns[] = {16, 64, 256};
        for (int mode = 0; mode < 2; mode++)
        for (int astep = 0; astep <= 256; astep += 256)
        for (int ni = 0; ni < 3; ni++)
            for (int reps = 1024; reps <= 1024; reps *= 4) {
                printf("mode=%d astep=%d  ", mode, astep);
                timing<<<1, 128, 65536>>>(reps, ns[ni], dc, astep, mode, 1);
                cudaError_t e2 = cudaDeviceSynchronize();
                long long hc[2]; cudaMemcpy(hc, dc, 16, cudaMemcpyDeviceToHost);
                printf("N=%3d reps=%4d err=%d: issue %.1f cycles/MMA, complete %.1f cycles/MMA\n", ns[ni], reps, (int)e2,
                       (double)hc[0] / reps, (double)hc[1] / reps);
            }
        for (int iss = 1; iss <= 4; iss++) {
            timing<<<1, 128, 65536>>>(1024, 16, dc, 0, 0, iss);
            cudaError_t e2 = cudaDeviceSynchronize();
            long long hc[2]; cudaMemcpy(hc, dc, 16, cudaMemcpyDeviceToHost);
            printf("issuers=%d err=%d: %.1f cycles per MMA overall (each issues 1024)\n", iss, (int)e2, (double)hc[0] / (1024.0 * iss));
        }
    }
    return 0;
}
