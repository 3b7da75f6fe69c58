// Probe: a 2-D TMA tensor map over a flat byte buffer with OVERLAPPING rows (inner extent 256, row
// stride 128) fetches 128 * LPR contiguous bytes from an arbitrary (unaligned) byte offset into
// 128-byte-aligned shared memory with the 128-byte swizzle.  Checks, for several offsets,
//   smem[128 * r + 16 * (k ^ ((abs_line(r)) & 7)) + i] == g[off + 128 r + 16 k + i]
// where abs_line = (smem byte address >> 7).  Also reports what out-of-bounds lines read as.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe_tma tools/probe_tma.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>

constexpr int LPR = 9;   // lines (128 B) per fetched row

__global__ void probe(const __grid_constant__ CUtensorMap tm, int c0, int c1, int dst_lines, uint8_t* out, uint32_t* sbase_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t sb = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t bb = (uint32_t)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bb));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) smem[i] = 0xEE;
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.shared::cta;");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bb), "r"(128 * LPR));
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(sb + 128 * dst_lines), "l"(&tm), "r"(c0), "r"(c1), "r"(bb) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bb));
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) out[i] = smem[i];
    if (threadIdx.x == 0) *sbase_out = sb;
}

int main(int argc, char** argv) {
    const int only = argc > 1 ? atoi(argv[1]) : -1;
    const int swz = argc > 2 ? atoi(argv[2]) : 3;
    const size_t total = 22 * 1125 * 7;   // 7 trials of 22 x 1125
    uint8_t* h = (uint8_t*)malloc(total);
    for (size_t i = 0; i < total; i++) h[i] = (uint8_t)((i * 131u + (i >> 8) * 7u + 5u) & 0xff);
    uint8_t *g, *out; uint32_t* sbo;
    cudaMalloc(&g, total); cudaMalloc(&out, 4096); cudaMalloc(&sbo, 4);
    cudaMemcpy(g, h, total, cudaMemcpyHostToDevice);
    typedef CUresult (*enc_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr);
    printf("entry point err=%d qr=%d fn=%p\n", (int)e, (int)qr, fn);
    if (!fn) return 1;
    CUtensorMap tm;
    const cuuint64_t nlines = (total - 256) / 128 + 1;       // every in-bounds element lies inside the buffer
    cuuint64_t dims[2] = {256, nlines};
    cuuint64_t strides[1] = {128};
    cuuint32_t box[2] = {128, LPR};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = ((enc_t)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, g, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             (CUtensorMapSwizzle)swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode result=%d nlines=%llu\n", (int)r, (unsigned long long)nlines);
    if (r != CUDA_SUCCESS) return 1;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192);
    uint8_t ho[4096];
    const size_t offs[] = {0, 0, 16, 5, 1125, 2 * 1125 + 24750, 24750 * 3 + 21 * 1125, 5, 127, 24750 * 5 + 7 * 1125, total - 1152, total - 1125, total - 1125 * 2};
    const int dls[] = {0, 9, 0, 0, 9, 18, 1, 3, 5, 9, 0, 9, 18};
    for (int t = 0; t < 13; t++) {
        if (only >= 0 && t != only) continue;
        const size_t off = offs[t];
        const int dl = dls[t];
        probe<<<1, 128, 8192>>>(tm, (int)(off % 128), (int)(off / 128), dl, out, sbo);
        e = cudaDeviceSynchronize();
        uint32_t sb;
        cudaMemcpy(ho, out, 4096, cudaMemcpyDeviceToHost);
        cudaMemcpy(&sb, sbo, 4, cudaMemcpyDeviceToHost);
        int bad = 0, zero_tail = 0, first_bad = -1;
        for (int rr = 0; rr < LPR; rr++)
            for (int k = 0; k < 8; k++)
                for (int i = 0; i < 16; i++) {
                    const size_t gi = off + 128 * rr + 16 * k + i;
                    const uint32_t line_abs = (sb >> 7) + dl + rr;
                    const int got = ho[128 * (dl + rr) + 16 * (k ^ (swz == 3 ? (line_abs & 7) : 0)) + i];
                    const bool inb = (off / 128 + rr) < nlines;   // element in bounds per the tensor map
                    if (!inb || gi >= total) { zero_tail += (got == 0); continue; }
                    if (got != h[gi]) { bad++; if (first_bad < 0) first_bad = 128 * rr + 16 * k + i; }
                }
        printf("off=%zu (c0=%d c1=%d) dst_line=%d sbase=0x%x err=%d mismatches=%d first_bad=%d oob_zero_bytes=%d untouched_before=%d\n", off,
               (int)(off % 128), (int)(off / 128), dl, sb, (int)e, bad, first_bad, zero_tail, dl > 0 ? (ho[128 * dl - 1] == 0xEE) : -1);
    }
    return 0;
}
