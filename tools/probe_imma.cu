// Probe: issue rate and latency of mma.sync m16n8k32.s8 (IMMA.16832.S8.S8) and m16n8k16.s8 on sm_100a.
//   probe_imma            -> cycles per IMMA per scheduler for 1..8 warps per scheduler, with 1/2/4/8
//                            independent accumulator chains per warp (1 chain = dependent-issue latency)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe_imma tools/probe_imma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int CHAINS, int K32>
__global__ void k(int iters, int* out, long long* cyc) {
    int acc[CHAINS][4];
    uint32_t a[4] = {threadIdx.x * 3u, threadIdx.x * 5u, 7u, 11u}, b0 = threadIdx.x, b1 = 13u;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = c;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CHAINS; c++) {
            if (K32)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(acc[c][0]), "+r"(acc[c][1]), "+r"(acc[c][2]), "+r"(acc[c][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
            else
                asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                             : "+r"(acc[c][0]), "+r"(acc[c][1]), "+r"(acc[c][2]), "+r"(acc[c][3])
                             : "r"(a[0]), "r"(a[1]), "r"(b0));
        }
    }
    const long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) s += acc[c][0] + acc[c][1] + acc[c][2] + acc[c][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int CHAINS, int K32>
void run(int warps_per_sched, int* out, long long* cyc) {
    const int iters = 2000, threads = 128 * warps_per_sched;
    k<CHAINS, K32><<<148, threads>>>(iters, out, cyc);
    k<CHAINS, K32><<<148, threads>>>(iters, out, cyc);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < 148; i++) avg += h[i];
    avg /= 148;
    // IMMAs per scheduler = warps_per_sched * CHAINS * iters
    printf("  %s chains=%d warps/sched=%d : %.2f cycles per IMMA per scheduler (%.1f cycles per loop trip per warp)\n",
           K32 ? "k32" : "k16", CHAINS, warps_per_sched, avg / (double)(warps_per_sched * CHAINS * iters), avg / iters);
}

int main() {
    int* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    for (int w = 1; w <= 8; w *= 2) {
        run<1, 1>(w, out, cyc); run<2, 1>(w, out, cyc); run<4, 1>(w, out, cyc); run<8, 1>(w, out, cyc);
    }
    for (int w = 1; w <= 8; w *= 2) { run<1, 0>(w, out, cyc); run<4, 0>(w, out, cyc); }
    printf("err=%d\n", (int)cudaGetLastError());
    return 0;
}
