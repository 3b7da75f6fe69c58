// Probe: stmatrix .m16n8 .trans .b8 on sm_100a -- where does byte i of lane l's register land?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void probe(uint8_t* out) {
    __shared__ __align__(128) uint8_t st[4][8][16];
    int lane = threadIdx.x;
    for (int i = lane; i < 512; i += 32) ((uint8_t*)st)[i] = 0xff;
    __syncwarp();
    uint32_t r[4];
    for (int j = 0; j < 4; j++) {   // byte i of lane l = 4*l + i  (0..127), matrix j tagged in bit 7 for j odd only -> use separate runs
        r[j] = (uint32_t)(4 * lane) * 0x01010101u + 0x03020100u;
    }
    uint32_t saddr = (uint32_t)__cvta_generic_to_shared(&st[lane >> 3][lane & 7][0]);
    asm volatile("stmatrix.sync.aligned.m16n8.x4.trans.shared.b8 [%0], {%1,%2,%3,%4};" ::"r"(saddr), "r"(r[0]), "r"(r[1]), "r"(r[2] | 0x80808080u), "r"(r[3] | 0x80808080u) : "memory");
    __syncwarp();
    for (int i = lane; i < 512; i += 32) out[i] = ((uint8_t*)st)[i];
}
int main() {
    uint8_t* d; cudaMalloc(&d, 512);
    probe<<<1, 32>>>(d);
    uint8_t h[512];
    cudaError_t e = cudaMemcpy(h, d, 512, cudaMemcpyDeviceToHost);
    printf("err=%d (%s)\n", (int)e, cudaGetErrorString(e));
    for (int m = 0; m < 4; m++)
        for (int r = 0; r < 8; r++) {
            printf("matrix %d row %d:", m, r);
            for (int c = 0; c < 16; c++) { int v = h[(m * 8 + r) * 16 + c]; printf(" %s%d.%d", (v & 0x80) ? "H" : "", (v & 0x7f) >> 2, v & 3); }
            printf("\n");
        }
    return 0;
}
