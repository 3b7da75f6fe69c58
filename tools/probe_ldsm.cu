// Probe: fragment layout of ldmatrix .m16n16 .trans .b8 on sm_100a (no PTX docs offline).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void probe(uint32_t* out) {
    __shared__ __align__(128) uint8_t m[2][16][16];
    for (int i = threadIdx.x; i < 512; i += 32) ((uint8_t*)m)[i] = (uint8_t)i;   // value = row*16+col (+256 for matrix 1 wraps)
    __syncwarp();
    int lane = threadIdx.x;
    uint32_t a0, a1, b0, b1, b2, b3;
    uint32_t addr = (uint32_t)__cvta_generic_to_shared(&m[0][lane & 15][0]);
    asm volatile("ldmatrix.sync.aligned.m16n16.x1.trans.shared.b8 {%0,%1}, [%2];" : "=r"(a0), "=r"(a1) : "r"(addr));
    uint32_t addr2 = (uint32_t)__cvta_generic_to_shared(&m[lane >> 4][lane & 15][0]);
    asm volatile("ldmatrix.sync.aligned.m16n16.x2.trans.shared.b8 {%0,%1,%2,%3}, [%4];" : "=r"(b0), "=r"(b1), "=r"(b2), "=r"(b3) : "r"(addr2));
    out[lane * 6 + 0] = a0; out[lane * 6 + 1] = a1;
    out[lane * 6 + 2] = b0; out[lane * 6 + 3] = b1; out[lane * 6 + 4] = b2; out[lane * 6 + 5] = b3;
}
int main() {
    uint32_t* d; cudaMalloc(&d, 32 * 6 * 4);
    probe<<<1, 32>>>(d);
    uint32_t h[32 * 6];
    cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("err=%d\n", (int)e);
    for (int l = 0; l < 32; l++) {
        printf("lane %2d:", l);
        for (int r = 0; r < 6; r++) {
            printf("  r%d=", r);
            for (int b = 0; b < 4; b++) { int v = (h[l * 6 + r] >> (8 * b)) & 255; printf("(%d,%d)", v >> 4, v & 15); }
        }
        printf("\n");
    }
    return 0;
}
