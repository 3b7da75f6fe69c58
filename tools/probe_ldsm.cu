// Probe: fragment layouts on sm_100a that the fused kernel relies on (no PTX docs offline):
//   ldmatrix .m16n16 .trans .b8 (x1 / x2 register order), stmatrix .m8n8 .x4 .b16, mad.hi.s32.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void probe(uint32_t* out) {
    __shared__ __align__(128) uint8_t m[2][16][16];
    __shared__ __align__(128) uint8_t st[4][8][16];
    for (int i = threadIdx.x; i < 512; i += 32) ((uint8_t*)m)[i] = i < 256 ? (uint8_t)i : (uint8_t)(((i & 15) << 4) | ((i >> 4) & 15));   // matrix 1 holds (col,row)
    __syncwarp();
    int lane = threadIdx.x;
    uint32_t a0, a1, b0, b1, b2, b3;
    uint32_t addr = (uint32_t)__cvta_generic_to_shared(&m[0][lane & 15][0]);
    asm volatile("ldmatrix.sync.aligned.m16n16.x1.trans.shared.b8 {%0,%1}, [%2];" : "=r"(a0), "=r"(a1) : "r"(addr));
    uint32_t addr2 = (uint32_t)__cvta_generic_to_shared(&m[lane >> 4][lane & 15][0]);
    asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];" : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3) : "r"(addr2));
    out[lane * 8 + 0] = a0; out[lane * 8 + 1] = a1;
    out[lane * 8 + 2] = b0; out[lane * 8 + 3] = b1; out[lane * 8 + 4] = b2; out[lane * 8 + 5] = b3;
    // stmatrix: register j of lane l = 0x(j)(l) pattern; read back which smem bytes it lands in
    uint32_t saddr = (uint32_t)__cvta_generic_to_shared(&st[lane >> 3][lane & 7][0]);
    uint32_t r[4];
    for (int j = 0; j < 4; j++) r[j] = (uint32_t)(lane | (j << 5)) * 0x01010101u + 0x03020100u * 0;   // 4 equal bytes
    asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1,%2,%3,%4};" ::"r"(saddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
    __syncwarp();
    // each lane reports the 4 bytes at st[lane>>3][lane&7][0..3] and [4..7]
    out[lane * 8 + 6] = *reinterpret_cast<uint32_t*>(&st[lane >> 3][lane & 7][0]) & 0xff
                      | (*reinterpret_cast<uint32_t*>(&st[lane >> 3][lane & 7][4]) & 0xff) << 8
                      | (*reinterpret_cast<uint32_t*>(&st[lane >> 3][lane & 7][8]) & 0xff) << 16
                      | (*reinterpret_cast<uint32_t*>(&st[lane >> 3][lane & 7][12]) & 0xff) << 24;
    int hi;
    asm("mad.hi.s32 %0, %1, %2, %3;" : "=r"(hi) : "r"(-1000000 - lane), "r"(4000000), "r"(7));
    out[lane * 8 + 7] = (uint32_t)hi;
}
int main() {
    uint32_t* d; cudaMalloc(&d, 32 * 8 * 4);
    probe<<<1, 32>>>(d);
    uint32_t h[32 * 8];
    cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("err=%d\n", (int)e);
    for (int l = 0; l < 32; l++) {
        printf("lane %2d:", l);
        for (int r = 0; r < 6; r++) {
            printf("  r%d=", r);
            for (int b = 0; b < 4; b++) { int v = (h[l * 8 + r] >> (8 * b)) & 255; printf("(%d,%d)", v >> 4, v & 15); }
        }
        printf("  stm(row l&7 of matrix l>>3, bytes 0,4,8,12 hold reg|lane)=");
        for (int b = 0; b < 4; b++) { int v = (h[l * 8 + 6] >> (8 * b)) & 255; printf("[j%d l%d]", v >> 5, v & 31); }
        long long ref = ((long long)(-1000000 - l) * 4000000LL) >> 32;
        printf("  madhi=%d expect %lld\n", (int)h[l * 8 + 7], ref + 7);
    }
    return 0;
}
