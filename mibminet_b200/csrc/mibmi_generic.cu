// Generic reference-order kernel: any (C, T, F2, N), both input layouts, both reorder_bn modes,
// hardware integer division, optional per-layer taps.  One CTA per trial (grid-stride), all
// inter-layer activations in shared memory.  This is the exact/slow path of the library and the
// fallback for networks the fused kernel does not cover; it is also what forward_int8_taps runs.
//
// Follows edge-eegnet_wolf/src/cl/net/layer1.c:87-96, layer2.c:87-118 / :139-210, layer3.c:70,
// layer4.c:99-140 (plain variant without per-element clip, :115-117), layer5.c:81-89.
#include "mibmi_common.cuh"
#include "mibmi_kernels.h"

namespace mibmi {

__global__ void __launch_bounds__(256)
generic_forward_kernel(const GenericNet* __restrict__ net, const int8_t* __restrict__ x, long long B,
                       int layout, int8_t* __restrict__ logits, int8_t* __restrict__ tap1,
                       int8_t* __restrict__ tap2, int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(16) int8_t smem[];
    const int C = net->C, T = net->T, F2 = net->F2, N = net->N, T8 = net->T8, T64 = net->T64;
    const bool reorder = net->reorder != 0;
    const int CA = (C + 3) / 4 * 4;
    const int y1s = T + 63, y2s = T8 + 15;
    int8_t* y1 = smem;                       // [F2][T+63]   31 zeros | T | 32 zeros
    int8_t* y2 = y1 + (size_t)F2 * y1s;      // [F2][T8+15]   7 zeros | T8 | 8 zeros
    int8_t* y3 = y2 + (size_t)F2 * y2s;      // [F2][T8]
    int8_t* y4 = y3 + (size_t)F2 * T8;       // [F2][T64]
    const int tid = threadIdx.x, nth = blockDim.x;

    for (long long b = blockIdx.x; b < B; b += gridDim.x) {
        const int8_t* xb = x + (size_t)b * (layout == 0 ? (size_t)C * T : (size_t)T * CA);
        for (int i = tid; i < F2 * y1s + F2 * y2s; i += nth) smem[i] = 0;
        __syncthreads();
        // ---- layer 1: spatial conv + BN
        for (int i = tid; i < F2 * T; i += nth) {
            const int k = i / T, t = i - k * T;
            const int8_t* w = net->w1 + k * C;
            int a = 0;
            if (layout == 0) for (int c = 0; c < C; c++) a += (int)xb[(size_t)c * T + t] * (int)w[c];
            else             for (int c = 0; c < C; c++) a += (int)xb[(size_t)t * CA + c] * (int)w[c];
            const int v = clip8((a + net->off1[k]) / net->fac1[k]);
            y1[k * y1s + 31 + t] = (int8_t)v;
            if (tap1) tap1[((size_t)b * F2 + k) * T + t] = (int8_t)v;
        }
        __syncthreads();
        // ---- layer 2: depthwise temporal conv (64 taps) + BN + ReLU + pool 8
        for (int i = tid; i < F2 * T8; i += nth) {
            const int k = i / T8, q = i - k * T8;
            const int8_t* w = net->w2 + k * kNf;
            const int8_t* r = y1 + k * y1s + q * 8;
            const int off = net->off2[k], fac = net->fac2[k];
            int s = 0;
            for (int e = 0; e < 8; e++) {
                int v = 0;
                for (int j = 0; j < kNf; j++) v += (int)r[e + j] * (int)w[j];
                if (reorder) s += max(v, -(off >> 3));
                else         s += max(clip8((v + (off >> 3)) / (fac >> 3)), 0);
            }
            const int o = reorder ? clip8((s + off) / fac) : clip8(s >> 3);
            y2[k * y2s + 7 + q] = (int8_t)o;
            if (tap2) tap2[((size_t)b * F2 + k) * T8 + q] = (int8_t)o;
        }
        __syncthreads();
        // ---- layer 3: depthwise temporal conv (16 taps), scalar requant
        for (int i = tid; i < F2 * T8; i += nth) {
            const int k = i / T8, t = i - k * T8;
            const int8_t* w = net->w3 + k * kL3Taps;
            const int8_t* r = y2 + k * y2s + t;
            int a = 0;
            for (int j = 0; j < kL3Taps; j++) a += (int)r[j] * (int)w[j];
            const int o = clip8(a / net->fac3);
            y3[k * T8 + t] = (int8_t)o;
            if (tap3) tap3[((size_t)b * F2 + k) * T8 + t] = (int8_t)o;
        }
        __syncthreads();
        // ---- layer 4: pointwise conv + BN + ReLU + pool 8
        for (int i = tid; i < F2 * T64; i += nth) {
            const int k = i / T64, q = i - k * T64;
            const int8_t* w = net->w4 + k * F2;
            const int off = net->off4[k], fac = net->fac4[k];
            int s = 0;
            for (int e = 0; e < 8; e++) {
                int v = 0;
                for (int c = 0; c < F2; c++) v += (int)y3[c * T8 + q * 8 + e] * (int)w[c];
                if (reorder) s += max(v, -(off >> 3));
                else         s += max((v + (off >> 3)) / (fac >> 3), 0);   // no clip: layer4.c:115-117
            }
            const int o = reorder ? clip8((s + off) / fac) : clip8(s >> 3);
            y4[k * T64 + q] = (int8_t)o;
            if (tap4) tap4[((size_t)b * F2 + k) * T64 + q] = (int8_t)o;
        }
        __syncthreads();
        // ---- layer 5: classifier
        for (int n = tid; n < N; n += nth) {
            const int8_t* w = net->w5 + (size_t)n * F2 * T64;
            int z = net->b5[n];
            for (int i = 0; i < F2 * T64; i++) z += (int)y4[i] * (int)w[i];
            logits[(size_t)b * N + n] = (int8_t)clip8(z / net->fac5);
        }
        __syncthreads();
    }
}

size_t generic_smem_bytes(int T, int F2) {
    const int T8 = T / 8, T64 = T8 / 8;
    return (size_t)F2 * ((T + 63) + (T8 + 15) + T8 + T64) + 16;
}

cudaError_t launch_generic(const GenericNet* net_dev, int T, int F2, const int8_t* x, long long B, int layout,
                           int8_t* logits, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                           cudaStream_t stream) {
    const size_t smem = generic_smem_bytes(T, F2);
    cudaError_t e = cudaFuncSetAttribute(generic_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    long long grid = (long long)num_sms * 4;
    if (grid > B) grid = B;
    generic_forward_kernel<<<(unsigned)grid, 256, smem, stream>>>(net_dev, x, B, layout, logits, t1, t2, t3, t4);
    return cudaGetLastError();
}

// ---- pre / post ops ------------------------------------------------------------------------------
// gen_input_header.py:73 / functional.py:308-335: trunc(clip(x / s, -1, 1) * 127) in float32.
__device__ __forceinline__ int quantize_one(float x, float s) {
    float v = __fdiv_rn(x, s);                       // IEEE division: the result must match NumPy's float32 x / s bit for bit
    v = fminf(fmaxf(v, -1.0f), 1.0f);
    v = __fdiv_rn(__fmul_rn(v, 254.0f), 2.0f);
    return (int)v;                                   // float -> int conversion truncates toward zero
}
__global__ void quantize_input_kernel(const float* __restrict__ x, long long n, float s, int8_t* __restrict__ q) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        q[i] = (int8_t)quantize_one(x[i], s);
}
// 16 floats per thread: four 16-byte loads in flight, one 16-byte store (x and q 16-byte aligned, n16 = n / 16 groups)
__global__ void __launch_bounds__(256) quantize_input_kernel_v16(const float4* __restrict__ x, long long n16, float s,
                                                                 uint4* __restrict__ q) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += (long long)gridDim.x * blockDim.x) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = __ldcs(x + 4 * i + k);          // streamed once
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            w[k] = (uint32_t)(quantize_one(v[k].x, s) & 255) | ((uint32_t)(quantize_one(v[k].y, s) & 255) << 8) |
                   ((uint32_t)(quantize_one(v[k].z, s) & 255) << 16) | ((uint32_t)(quantize_one(v[k].w, s) & 255) << 24);
        q[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}
// functional.py:338-353: x / 127 * s  (float64 in the reference; float32 here, documented tolerance)
__global__ void dequantize_kernel(const int8_t* __restrict__ q, long long n, float s, float* __restrict__ y) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        y[i] = (float)((double)q[i] / 127.0 * (double)s);
}

cudaError_t launch_quantize(const float* x, long long n, float s, int8_t* q, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    long long body = 0;
    if ((((uintptr_t)x | (uintptr_t)q) & 15) == 0 && n >= 16) {
        body = n & ~15ll;
        long long blocks = (body / 16 + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;
        quantize_input_kernel_v16<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(x), body / 16, s,
                                                                         reinterpret_cast<uint4*>(q));
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess || body == n) return e;
    }
    long long blocks = (n - body + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    quantize_input_kernel<<<(unsigned)blocks, 256, 0, stream>>>(x + body, n - body, s, q + body);
    return cudaGetLastError();
}
cudaError_t launch_dequantize(const int8_t* q, long long n, float s, float* y, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    dequantize_kernel<<<(unsigned)blocks, 256, 0, stream>>>(q, n, s, y);
    return cudaGetLastError();
}

}  // namespace mibmi
