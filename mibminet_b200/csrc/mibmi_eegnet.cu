// EEGNet / Q-EEGNet layer order: temporal conv (1, Nf) over every EEG channel FIRST, then the depthwise
// spatial conv -- QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:100-106 (modules), :137-168 (forward).  The reference
// tree holds no integer model of this topology; the integer semantics are the edgeEEGNet ones (golden_model.py:153-379:
// truncating division, clip to [-128, 127], REORDER_BN pooling) applied to this layer order -- oracle/eegnet_np.py.
//
//   L1  v[f][c][t] = sum_j xpad[c][t + j] w1[f][j]   (Nf/2 - 1 zeros | x | Nf/2 zeros);  y1 = clip8(tdiv(v + off1[f], fac1[f]))
//   L2  a[k][t] = sum_c y1[k / D][c][t] w2[k][c];  s = sum_8 max(a, -(off2[k] >> 3));  y2 = clip8(tdiv(s + off2[k], fac2[k]))
//   L3..L5 as in mibmi_generic.cu.
//
// One CTA per trial; layers 1 and 2 are fused so that the F1 x C x T intermediate (198 kB at 22 x 1125) never exists.
//   eegnet_mma_kernel      (default, Nf a multiple of 32): warp = temporal filter, the Nf-tap convolution as a Hankel x Toeplitz
//                          product on mma.sync.m16n8k32.s8 (the dense contraction of this topology), requant + spatial conv +
//                          pooling in the epilogue -- see the comment above the kernel.
//   eegnet_forward_kernel  (fallback, any Nf <= 128 that is a multiple of 4): a thread owns 4 consecutive samples of 4 filters,
//                          walks over the EEG channels, convolves with dp4a on ALIGNED words of the padded row against 4
//                          byte-shifted copies of the taps (output t0 + s = sum_m xpad[t0 + m] W_s[m], W_s[m] = w[m - s]).
// Both requantise with the magic division of DESIGN.md section 5 when every layer-1 factor is inside its proven range, and
// with the hardware division otherwise (exact for any factor).
#include <atomic>
#include <new>
#include <stdlib.h>
#include <string.h>

#include "../../include/mibmi.h"
#include "mibmi_common.cuh"
#include "mibmi_ptx.cuh"

namespace mibmi {

#ifndef MIBMI_EEG_CUNROLL
#define MIBMI_EEG_CUNROLL 1
#endif
constexpr int kEegCUnroll = MIBMI_EEG_CUNROLL;       // channel loop of the tensor-core kernel
constexpr int kEegMaxF1 = 16, kEegMaxK2 = 32, kEegMaxNf = 128, kEegMaxNW = kEegMaxNf / 4 + 1;

struct EegNet {
    int C, T, F1, D, K2, F2, N, Nf, T8, T64, XP, G, NW;
    int fac1[kEegMaxF1], off1[kEegMaxF1];
    int mg1[kEegMaxF1], sh1[kEegMaxF1], sg1[kEegMaxF1];    // layer-1 magic division (valid when `magic1`): m, shift, sign of the factor
    int magic1;
    int fac2[kEegMaxK2], off2[kEegMaxK2];
    int fac4[kMaxF2], off4[kMaxF2];
    int fac3, fac5;
    int b5[kMaxN];
    alignas(16) int w1s[kEegMaxF1 * kEegMaxNW * 4];   // [f][word i][shift s]
    int8_t w2[kEegMaxK2 * kMaxC];                      // [K2][C]
    int8_t w3[kEegMaxK2 * kL3Taps];                    // [K2][16] torch order
    int8_t w4[kMaxF2 * kEegMaxK2];                     // [F2][K2]
    const int8_t* w5;                                  // [N][F2 * T64] device pointer
    const uint32_t* w1frag;                            // tensor-core kernel: Toeplitz B fragments [F1][4 v][KS][2 h][32 lanes], device pointer
    int KS, MT, XPM;                                   // k-steps of 32 (= (Nf + 63) / 32), m-tiles of 16 rows x 32 outputs, row pitch
    // tcgen05 kernel (eegnet_tc_kernel): chunks of 128 samples, row pitch of a padded channel, tiles of 16 channels, input stages
    int NCH, P, CT, NXS;
    int m32[kEegMaxF1];                                // layer 1, "no shift" magic (floor(2^32 / |fac|) + 1), factor sign folded into taps / offset
    long long offm32[kEegMaxF1];                       // sg * off1 * m32: hi32(x m + off m) = hi32((x + off) m)
    int magic_rest;                                    // 1: layers 2 - 5 requantise with the magic division below (else hardware division)
    int mg2[kEegMaxK2], sh2[kEegMaxK2], sg2[kEegMaxK2], mg4[kMaxF2], sh4[kMaxF2], sg4[kMaxF2], mg3, sh3, sg3, mg5, sh5, sg5;
    const uint8_t* b1umma;                             // Toeplitz operand of the temporal conv [128 = (f, n)][160] as K-major core matrices
    const uint8_t* b2umma;                             // block-diagonal spatial weights [16 = k][8 * 16 CT = (f, c)] as K-major core matrices
};

__device__ __forceinline__ int eclip8(int v) { return min(max(v, -128), 127); }
__device__ __forceinline__ int esat8(int v) {          // the same clip as one saturating conversion
    int r;
    asm("cvt.sat.s8.s32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}

// layers 3, 4, 5 of the trial whose pooled layer-2 output sits in y2 (zero padded rows); all threads of the CTA
__device__ __forceinline__ void eeg_layers345(const EegNet* __restrict__ net, const int8_t* y2, int8_t* y3, int8_t* y4, long long b,
                                              int8_t* __restrict__ logits, int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    const int K2 = net->K2, F2 = net->F2, N = net->N, T8 = net->T8, T64 = net->T64, y2s = T8 + 15;
    const int tid = threadIdx.x, nth = blockDim.x;
    // ---- layer 3 (eegnet.py:155-156)
    for (int i = tid; i < K2 * T8; i += nth) {
        const int k = i / T8, t = i - k * T8;
        const int8_t* w = net->w3 + k * kL3Taps;
        const int8_t* r = y2 + k * y2s + t;
        int acc = 0;
        for (int j = 0; j < kL3Taps; j++) acc += (int)r[j] * (int)w[j];
        const int o = eclip8(acc / net->fac3);
        y3[k * T8 + t] = (int8_t)o;
        if (tap3) tap3[((size_t)b * K2 + k) * T8 + t] = (int8_t)o;
    }
    __syncthreads();
    // ---- layer 4 (eegnet.py:157-161)
    for (int i = tid; i < F2 * T64; i += nth) {
        const int k = i / T64, q = i - k * T64;
        const int8_t* w = net->w4 + k * K2;
        const int off = net->off4[k], fac = net->fac4[k];
        int s = 0;
        for (int e = 0; e < 8; e++) {
            int v = 0;
            for (int c = 0; c < K2; c++) v += (int)y3[c * T8 + q * 8 + e] * (int)w[c];
            s += max(v, -(off >> 3));
        }
        const int o = eclip8((s + off) / fac);
        y4[k * T64 + q] = (int8_t)o;
        if (tap4) tap4[((size_t)b * F2 + k) * T64 + q] = (int8_t)o;
    }
    __syncthreads();
    // ---- layer 5 (eegnet.py:164-166)
    for (int n = tid >> 5; n < N; n += nth >> 5) {            // warp = class: lanes split the features, shuffle reduction
        const int8_t* w = net->w5 + (size_t)n * F2 * T64;
        int z = 0;
        for (int i = tid & 31; i < F2 * T64; i += 32) z += (int)y4[i] * (int)w[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
        if ((tid & 31) == 0) logits[(size_t)b * N + n] = (int8_t)eclip8((z + net->b5[n]) / net->fac5);
    }
}

template <int D, bool MAGIC>
__global__ void __launch_bounds__(384)
eegnet_forward_kernel(const EegNet* __restrict__ net, const int8_t* __restrict__ x, long long B, int8_t* __restrict__ logits,
                      int8_t* __restrict__ tap2, int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(16) int8_t smem[];
    const int C = net->C, T = net->T, F1 = net->F1, K2 = net->K2, Nf = net->Nf;
    const int T8 = net->T8, XP = net->XP, G = net->G, NW = net->NW;
    const int y2s = T8 + 15;
    int4* w1s = reinterpret_cast<int4*>(smem);                                  // [F1][NW] int4
    int8_t* xs = smem + (size_t)F1 * NW * 16;                                   // [C][XP]
    int* part = reinterpret_cast<int*>(xs + (size_t)C * XP);                    // [K2][G]  (XP % 4 == 0)
    int8_t* y2 = reinterpret_cast<int8_t*>(part + (size_t)K2 * G);              // [K2][T8 + 15]
    int8_t* y3 = y2 + (size_t)K2 * y2s;                                         // [K2][T8]
    int8_t* y4 = y3 + (size_t)K2 * T8;                                          // [F2][T64]
    const int tid = threadIdx.x, nth = blockDim.x;
    const int PL = Nf / 2 - 1;

    for (int i = tid; i < F1 * NW; i += nth) w1s[i] = reinterpret_cast<const int4*>(net->w1s)[i];
    for (int i = tid; i < C * XP / 4; i += nth) reinterpret_cast<int*>(xs)[i] = 0;
    for (int i = tid; i < K2 * y2s; i += nth) y2[i] = 0;
    __syncthreads();

    for (long long b = blockIdx.x; b < B; b += gridDim.x) {
        const int8_t* xb = x + (size_t)b * C * T;
        for (int i = tid; i < C * T; i += nth) {
            const int c = i / T, t = i - c * T;
            xs[c * XP + PL + t] = xb[i];
        }
        __syncthreads();
        // ---- layers 1 + 2 (eegnet.py:143-151)
        for (int item = tid; item < (F1 / 4) * G; item += nth) {
            const int fh = item / G, g = item - fh * G;
            int a[4][D][4];
            int fac[4], off[4], shf[4], sgn[4];      // MAGIC: fac = magic multiplier
#pragma unroll
            for (int fi = 0; fi < 4; fi++) {
                fac[fi] = MAGIC ? net->mg1[4 * fh + fi] : net->fac1[4 * fh + fi];
                shf[fi] = net->sh1[4 * fh + fi];
                sgn[fi] = net->sg1[4 * fh + fi];
                off[fi] = net->off1[4 * fh + fi];
#pragma unroll
                for (int d = 0; d < D; d++) a[fi][d][0] = a[fi][d][1] = a[fi][d][2] = a[fi][d][3] = 0;
            }
            const int4* wf = w1s + (size_t)4 * fh * NW;
            for (int c = 0; c < C; c++) {
                const int* xr = reinterpret_cast<const int*>(xs + (size_t)c * XP) + g;
                int acc[4][4];
#pragma unroll
                for (int fi = 0; fi < 4; fi++) acc[fi][0] = acc[fi][1] = acc[fi][2] = acc[fi][3] = 0;
                for (int i = 0; i < NW; i++) {
                    const int xw = xr[i];
#pragma unroll
                    for (int fi = 0; fi < 4; fi++) {
                        const int4 w = wf[fi * NW + i];
                        acc[fi][0] = __dp4a(xw, w.x, acc[fi][0]);
                        acc[fi][1] = __dp4a(xw, w.y, acc[fi][1]);
                        acc[fi][2] = __dp4a(xw, w.z, acc[fi][2]);
                        acc[fi][3] = __dp4a(xw, w.w, acc[fi][3]);
                    }
                }
#pragma unroll
                for (int fi = 0; fi < 4; fi++) {
                    int w2v[D];
#pragma unroll
                    for (int d = 0; d < D; d++) w2v[d] = (int)net->w2[((4 * fh + fi) * D + d) * C + c];
#pragma unroll
                    for (int s = 0; s < 4; s++) {
                        const int xv = acc[fi][s] + off[fi];
                        const int y = eclip8(MAGIC ? sgn[fi] * tdiv_magic(xv, fac[fi], shf[fi]) : xv / fac[fi]);
#pragma unroll
                        for (int d = 0; d < D; d++) a[fi][d][s] += y * w2v[d];
                    }
                }
            }
#pragma unroll
            for (int fi = 0; fi < 4; fi++)
#pragma unroll
                for (int d = 0; d < D; d++) {
                    const int k = (4 * fh + fi) * D + d;
                    const int thr = -(net->off2[k] >> 3);
                    part[k * G + g] = max(a[fi][d][0], thr) + max(a[fi][d][1], thr) + max(a[fi][d][2], thr) + max(a[fi][d][3], thr);
                }
        }
        __syncthreads();
        for (int i = tid; i < K2 * T8; i += nth) {
            const int k = i / T8, q = i - k * T8;
            const int s = part[k * G + 2 * q] + part[k * G + 2 * q + 1];
            const int o = eclip8((s + net->off2[k]) / net->fac2[k]);
            y2[k * y2s + 7 + q] = (int8_t)o;
            if (tap2) tap2[((size_t)b * K2 + k) * T8 + q] = (int8_t)o;
        }
        __syncthreads();
        eeg_layers345(net, y2, y3, y4, b, logits, tap3, tap4);
        __syncthreads();
    }
}


// ---------------------------------------------------------------------------------------------------------
#ifndef MIBMI_EEG_MINB
#define MIBMI_EEG_MINB 2      // CTAs of 256 threads per SM the register allocation aims at
#endif
// Tensor-core form of layers 1 + 2 (the temporal conv is the dense contraction of this topology: F1 x C x T x Nf MACs).
// Warp f = temporal filter f.  Per EEG channel the Nf-tap convolution is a Hankel x Toeplitz product on mma.sync.m16n8k32.s8:
//   A[16 rows x 32]  = the padded input row itself, row r = bytes 32 r .. (ldmatrix.x4, rows 32 B apart, k-step s adds 32 s)
//   B[32 x 8]        = Toeplitz of the filter's taps: B[kk][n] = w1[f][32 s + kk - (8 v + n)], 4 variants v = outputs 8 v .. 8 v + 7
// so D[r][8 v + n] is conv output t = 32 r + 8 v + n; KS = (Nf + 63) / 32 k-steps, 4 KS IMMA per 512 outputs.  The filter's
// 8 KS fragment words stay in registers for the whole launch.  The epilogue requantises (batch_norm1 / quant2), multiplies by
// the D spatial weights of this filter and accumulates over the channels in registers; a pooling group of 8 outputs is one
// (row, variant) = 4 lanes x 2 values, reduced with two shuffles, requantised (batch_norm2 + ReLU + pool, quant3) and stored.
template <int D, int KS, bool MAGIC, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT <= 256 ? MIBMI_EEG_MINB : 1)
eegnet_mma_kernel(const EegNet* __restrict__ net, const int8_t* __restrict__ x, long long B, int8_t* __restrict__ logits,
                  int8_t* __restrict__ tap2, int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(16) int8_t smem[];
    const int C = net->C, T = net->T, K2 = net->K2, Nf = net->Nf;
    const int T8 = net->T8, XP = net->XPM, MT = net->MT;
    const int y2s = T8 + 15;
    int8_t* xs = smem;                                                          // [C][XP], XP % 16 == 0
    int8_t* y2 = xs + (size_t)C * XP;                                           // [K2][T8 + 15]
    int8_t* y3 = y2 + (size_t)K2 * y2s;                                         // [K2][T8]
    int8_t* y4 = y3 + (size_t)K2 * T8;                                          // [F2][T64]
    const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, f = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const int PL = Nf / 2 - 1;

    uint32_t bf[4][KS][2];
#pragma unroll
    for (int v = 0; v < 4; v++)
#pragma unroll
        for (int s = 0; s < KS; s++)
#pragma unroll
            for (int h = 0; h < 2; h++) bf[v][s][h] = net->w1frag[(((f * 4 + v) * KS + s) * 2 + h) * 32 + lane];
    const int off1 = MAGIC ? net->sg1[f] * net->off1[f] : net->off1[f], fac1 = MAGIC ? net->mg1[f] : net->fac1[f], shf1 = net->sh1[f];
    int off2[D], fac2[D], thr2[D];
#pragma unroll
    for (int d = 0; d < D; d++) {
        off2[d] = net->off2[f * D + d];
        fac2[d] = net->fac2[f * D + d];
        thr2[d] = -(off2[d] >> 3);
    }
    for (int i = tid; i < C * XP / 4; i += nth) reinterpret_cast<int*>(xs)[i] = 0;
    for (int i = tid; i < K2 * y2s; i += nth) y2[i] = 0;
    __syncthreads();
    // ldmatrix.x4 lane addressing: matrix j = lane >> 3, row (lane & 7) + 8 (j & 1), k-half j >> 1
    const uint32_t a_off = 32 * ((lane & 7) + 8 * ((lane >> 3) & 1)) + 16 * (lane >> 4);

    for (long long b = blockIdx.x; b < B; b += gridDim.x) {
        const int8_t* xb = x + (size_t)b * C * T;
{
            // aligned 32-bit global loads of the (arbitrarily aligned) trial, bytes scattered to their padded rows; (c, t) advance
            // incrementally (no division per element)
            const uintptr_t ga = reinterpret_cast<uintptr_t>(xb);
            const int mis = (int)(ga & 3), total = C * T, nwords = (mis + total + 3) >> 2;
            const uint32_t* gw = reinterpret_cast<const uint32_t*>(ga - mis);
            int i = 4 * tid - mis;
            int c = i >= 0 ? i / T : 0, t = i - c * T;
            for (int w = tid; w < nwords; w += 4 * nth) {              // 4 independent loads in flight per thread
                uint32_t v4[4];
#pragma unroll
                for (int u = 0; u < 4; u++) v4[u] = w + u * nth < nwords ? __ldg(gw + w + u * nth) : 0u;
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const uint32_t v = v4[u];
                    int cj = c, tj = t;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (tj >= T) { tj -= T; cj++; }
                        if (i + j >= 0 && i + j < total) xs[cj * XP + PL + tj] = (int8_t)(v >> (8 * j));
                        tj++;
                    }
                    i += 4 * nth;
                    t += 4 * nth;
                    while (t >= T) { t -= T; c++; }
                }
            }
        }
        __syncthreads();
        for (int mt = 0; mt < MT; mt++) {
            int a[D][4][4];
#pragma unroll
            for (int d = 0; d < D; d++)
#pragma unroll
                for (int v = 0; v < 4; v++) a[d][v][0] = a[d][v][1] = a[d][v][2] = a[d][v][3] = 0;
            const uint32_t abase = smem_u32(xs) + 512 * mt + a_off;
#pragma unroll kEegCUnroll
            for (int c = 0; c < C; c++) {
                int acc[4][4];                                  // the BN offset rides in the accumulator
#pragma unroll
                for (int v = 0; v < 4; v++) acc[v][0] = acc[v][1] = acc[v][2] = acc[v][3] = off1;
#pragma unroll
                for (int s = 0; s < KS; s++) {
                    uint32_t af[4];
                    ldsm_x4(af, abase + c * XP + 32 * s);
#pragma unroll
                    for (int v = 0; v < 4; v++) mma_k32(acc[v], af, bf[v][s][0], bf[v][s][1]);
                }
                int w2v[D];
#pragma unroll
                for (int d = 0; d < D; d++) w2v[d] = (int)net->w2[(f * D + d) * C + c];
#pragma unroll
                for (int v = 0; v < 4; v++)
#pragma unroll
                    for (int ci = 0; ci < 4; ci++) {
                        const int xv = acc[v][ci];
                        const int y = esat8(MAGIC ? tdiv_magic(xv, fac1, shf1) : xv / fac1);      // MAGIC: factor sign folded into taps / offset
#pragma unroll
                        for (int d = 0; d < D; d++) a[d][v][ci] += y * w2v[d];
                    }
            }
            // pooling group (row, v): outputs 32 row + 8 v .. + 7 = the 4 lanes q of a row x (ci & 1)
#pragma unroll
            for (int d = 0; d < D; d++)
#pragma unroll
                for (int v = 0; v < 4; v++)
#pragma unroll
                    for (int hf = 0; hf < 2; hf++) {
                        int p = max(a[d][v][2 * hf], thr2[d]) + max(a[d][v][2 * hf + 1], thr2[d]);
                        p += __shfl_xor_sync(0xffffffffu, p, 1);
                        p += __shfl_xor_sync(0xffffffffu, p, 2);
                        const int pg = 4 * (16 * mt + g + 8 * hf) + v;          // pooled sample index
                        if (q == 0 && pg < T8) {
                            const int k = f * D + d;
                            const int o = eclip8((p + off2[d]) / fac2[d]);
                            y2[k * y2s + 7 + pg] = (int8_t)o;
                            if (tap2) tap2[((size_t)b * K2 + k) * T8 + pg] = (int8_t)o;
                        }
                    }
        }
        __syncthreads();
        eeg_layers345(net, y2, y3, y4, b, logits, tap3, tap4);
        __syncthreads();
    }
}


// ---------------------------------------------------------------------------------------------------------
// tcgen05 form of layers 1 + 2 for F1 = 8 temporal filters of Nf = 128 taps (the QuantLab EEGNet configuration, eegnet.py:100-106):
// BOTH contractions run on the 5th-generation tensor core with their accumulators in TMEM, the warps only requantise.
//   MMA 1 (temporal conv, eegnet.py:143-145): per chunk of 128 samples and tile of 16 EEG channels ONE accumulator tile
//         D1[128 rows = (channel, block of 16 samples)][128 columns = (filter f, sample n of the block)] = Hankel(x) * Toeplitz(w1):
//         the A operand is the padded input row itself, described with overlapping strides (row = bytes 16 r .., LBO = 16 B: the next
//         16 K positions are the next 16 bytes; SBO = row pitch: the next 8 rows are the next channel) -- no im2col; K = 160 in 5 k-steps.
//   E1    thread = TMEM lane = (channel, block): per filter 16 accumulators -> batch_norm1 / quant2 (hi32(x m + off m), sign, saturating
//         pack) -> ONE 16-byte store into the A operand of MMA 2: [row (f, c)][128 samples], 128-byte swizzle, MN-major (the layout a
//         TMA tile [C][128] has in mibmi_tc.cu)
//   MMA 2 (depthwise spatial conv, eegnet.py:147): D2[128 samples][16 = f D + d] = A2[128][(f, c)] * blockdiag(w2), K = 8 x 16 CT
//   E2    batch_norm2 + ReLU + sum-pool 8 (8 consecutive TMEM lanes = 8 lanes of the m16n8 fragment: three shuffles) + quant3 -> y2
//   L     layers 3 - 5 (eeg_layers345_sub) on the double-buffered y2, one trial behind
// Warps: L 0-3, E2 4-7, E1 8-19 (three sets of four TMEM lane quarters, one accumulator buffer each), S 20-23 (trial -> padded rows),
// M 24 (issues both MMAs).  One persistent CTA per SM; hand-off is mbarrier-only.
#ifndef MIBMI_ET_SLEEP_NS
#define MIBMI_ET_SLEEP_NS 0               // > 0: the roles with long waits (L, E2, S) poll with nanosleep instead of the suspending try_wait
                                          // (r2h9: 100 / 400 ns are slower, 12.06 / 12.20 vs 12.69 M trials/s)
#endif
#ifdef MIBMI_PROF
__device__ long long g_prof_et[148 * 32 * 4];
#define EWAIT(slot, ...) do { if (!mbar_test(__VA_ARGS__)) { const long long t0_ = clock64(); mbar_wait(__VA_ARGS__); pw_[slot] += clock64() - t0_; } } while (0)
#define EPROF_BEGIN long long pw_[3] = {0, 0, 0}; const long long pt0_ = clock64();
#define EPROF_END if ((threadIdx.x & 31) == 0) { long long* o_ = g_prof_et + (blockIdx.x * 32 + (threadIdx.x >> 5)) * 4; \
    o_[0] = clock64() - pt0_; o_[1] = pw_[0]; o_[2] = pw_[1]; o_[3] = pw_[2]; }
#else
#define EWAIT(slot, ...) mbar_wait(__VA_ARGS__)
#define EPROF_BEGIN
#define EPROF_END
#endif
#define EWAITS(slot, ...) do { if (MIBMI_ET_SLEEP_NS > 0) mbar_wait_sleep(__VA_ARGS__, MIBMI_ET_SLEEP_NS); else EWAIT(slot, __VA_ARGS__); } while (0)
#ifndef MIBMI_ET_SW
#define MIBMI_ET_SW 3                     // staging warps (r2h4: 3 instead of 4 = 24 warps, 80 registers: 9.96 -> 11.09 M trials/s at 22 x 1125)
#endif
#ifndef MIBMI_ET_M64
#define MIBMI_ET_M64 0                    // 1: a last channel tile of <= 8 channels (22 channels: 16 + 6) as an M = 64 instruction: bit-exact (the
                                          // TMEM lane of row i is 32 (i / 16) + i % 16), no faster (r2h8: 12.35 vs 12.72 M trials/s, the same 3 % that the
                                          // extra selects cost at 64 x 480, where no tile is short)
#endif
#ifndef MIBMI_ET_E1W
#define MIBMI_ET_E1W 12                   // layer-1 epilogue warps: sets of four (one TMEM accumulator buffer per set)
#endif
constexpr int kEtL = 4, kEtE2 = 4, kEtE1 = MIBMI_ET_E1W, kEtS = MIBMI_ET_SW, kEtSets = kEtE1 / 4;
constexpr int kEtWarpE2 = kEtL, kEtWarpE1 = kEtWarpE2 + kEtE2, kEtWarpS = kEtWarpE1 + kEtE1, kEtWarpM = kEtWarpS + kEtS;
constexpr int kEtThreads = (kEtWarpM + 1) * 32;
constexpr int kEtB1Bytes = 128 * 160, kEtTmT2 = 128 * kEtSets;
// barriers (8 bytes each)
constexpr int kEtXsFull = 0, kEtXsEmpty = 16, kEtT1Full = 32, kEtT1Empty = 56, kEtA2Full = 80, kEtA2Empty = 96, kEtT2Full = 112,
              kEtT2Empty = 128, kEtY2Full = 144, kEtY2Empty = 160, kEtTmemSlot = 176, kEtOffCons = 256, kEtOffB1 = 1024;
static_assert(kEtTmT2 + 32 <= 512, "TMEM map");

struct EtSmem { int off_b2, off_xs, off_a2, off_y2, off_y3, off_y4, total; };
__host__ __device__ inline int et_y2s(int T8) { return (T8 + 3) / 4 * 4 + 20; }     // y2 row: 7 zeros | T8 samples | zeros, whole words
__host__ __device__ inline EtSmem et_smem(int C, int T, int K2, int F2, int NCH, int P, int CT, int NXS) {
    EtSmem m;
    (void)C;
    const int T8 = T / 8, T64 = T8 / 8;
    m.off_b2 = kEtOffB1 + kEtB1Bytes;
    m.off_a2 = (m.off_b2 + 16 * 8 * 16 * CT + 1023) / 1024 * 1024;          // swizzle atoms are 1024-byte aligned
    m.off_xs = m.off_a2 + 2 * 8 * 16 * CT * 128;
    m.off_y2 = m.off_xs + NXS * 16 * CT * P;
    m.off_y3 = (m.off_y2 + 2 * K2 * et_y2s(T8) + 15) / 16 * 16;   // y3t [T8][16 channels], time-major
    m.off_y4 = m.off_y3 + 16 * T8;
    m.total = (m.off_y4 + F2 * T64 + 127) / 128 * 128;
    (void)NCH;
    return m;
}

__device__ __forceinline__ constexpr uint32_t et_idesc_i8_kk(int M, int N) {   // both operands K-major
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void et_tmem_ld_32x32b_x16(uint32_t (&r)[16], uint32_t taddr) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr));
}
__device__ __forceinline__ int et_div_fold(int x, int m, long long offm) {     // tdiv(x + off, d) up to saturation
    const int hi = (int)(((long long)x * m + offm) >> 32);
    return hi + (int)((unsigned)hi >> 31);
}
// tdiv(x, fac): magic form sg * tdiv_magic(x, m, sh) (|fac| >= 4, |x| < 2^29, proven per network at create) or the hardware division
__device__ __forceinline__ int et_div(int x, int fac, int m, int sh, int sg, int magic) {
    return magic ? sg * tdiv_magic(x, m, sh) : x / fac;
}
__device__ __forceinline__ void et_bar_sync_l() { asm volatile("bar.sync 1, %0;" ::"n"(kEtL * 32) : "memory"); }

// Layers 3, 4, 5 on the four L warps (128 threads, named barrier 1).  Eight threads per channel:
//   layer 3 (eegnet.py:155-156): thread (k, u) owns the outputs 4 u .. 4 u + 3 of channel k: 5 aligned words of the padded y2 row against
//            4 byte-shifted copies of the 16 taps (W_s[m] = w3[k][m - s]) = 20 dp4a; results stored time-major, y3t[t][16 channels]
//   layer 4 (eegnet.py:157-161): thread (k, e) owns sample e of every pooling group: one 16-byte load of y3t[t] against the 16 input
//            weights of output channel k = 4 dp4a, threshold, sum over the 8 threads of the group with three shuffles
//   layer 5 (eegnet.py:164-166): warp = class
struct EtLRegs { uint32_t w3[5][4]; uint32_t w4[4]; int thr4, off4, fac4, m4, sh4, sg4; };
__device__ __forceinline__ void et_l_setup(const EegNet* __restrict__ net, int tid, EtLRegs& R) {
    const int k = tid >> 3, K2 = net->K2, F2 = net->F2;
#pragma unroll
    for (int i = 0; i < 5; i++)
#pragma unroll
        for (int sft = 0; sft < 4; sft++) {
            uint32_t w = 0;
            for (int e = 0; e < 4; e++) {
                const int j = 4 * i + e - sft;
                const int v = (k < K2 && j >= 0 && j < kL3Taps) ? net->w3[k * kL3Taps + j] : 0;
                w |= (uint32_t)(uint8_t)(int8_t)v << (8 * e);
            }
            R.w3[i][sft] = w;
        }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        uint32_t w = 0;
        for (int e = 0; e < 4; e++) {
            const int c = 4 * i + e;
            const int v = (k < F2 && c < K2) ? net->w4[k * K2 + c] : 0;
            w |= (uint32_t)(uint8_t)(int8_t)v << (8 * e);
        }
        R.w4[i] = w;
    }
    R.off4 = k < F2 ? net->off4[k] : 0;
    R.fac4 = k < F2 ? net->fac4[k] : 1;
    R.thr4 = -(R.off4 >> 3);
    R.m4 = k < F2 ? net->mg4[k] : 0; R.sh4 = k < F2 ? net->sh4[k] : 0; R.sg4 = k < F2 ? net->sg4[k] : 1;
}
__device__ __forceinline__ void et_layers345(const EegNet* __restrict__ net, const EtLRegs& R, const int8_t* y2, int y2s, int8_t* y3t,
                                             int8_t* y4, long long b, int8_t* __restrict__ logits, int8_t* __restrict__ tap3,
                                             int8_t* __restrict__ tap4, int tid) {
    const int K2 = net->K2, F2 = net->F2, N = net->N, T8 = net->T8, T64 = net->T64;
    const int k = tid >> 3, e8 = tid & 7, fac3 = net->fac3, mg = net->magic_rest;
    const int m3 = net->mg3, sh3 = net->sh3, sg3 = net->sg3;
    if (k < K2) {
        const uint32_t* row = reinterpret_cast<const uint32_t*>(y2 + k * y2s);
        for (int u = e8; 4 * u < T8; u += 8) {
            int acc[4] = {0, 0, 0, 0};
#pragma unroll
            for (int i = 0; i < 5; i++) {
                const int xw = (int)row[u + i];
#pragma unroll
                for (int sft = 0; sft < 4; sft++) acc[sft] = __dp4a(xw, (int)R.w3[i][sft], acc[sft]);
            }
#pragma unroll
            for (int sft = 0; sft < 4; sft++) {
                const int t = 4 * u + sft;
                if (t < T8) {
                    const int o = eclip8(et_div(acc[sft], fac3, m3, sh3, sg3, mg));
                    y3t[t * 16 + k] = (int8_t)o;
                    if (tap3) tap3[((size_t)b * K2 + k) * T8 + t] = (int8_t)o;
                }
            }
        }
    }
    et_bar_sync_l();
    for (int q = 0; q < T64; q++) {
        const uint4 yv = *reinterpret_cast<const uint4*>(y3t + (8 * q + e8) * 16);
        int v = __dp4a((int)yv.x, (int)R.w4[0], 0);
        v = __dp4a((int)yv.y, (int)R.w4[1], v);
        v = __dp4a((int)yv.z, (int)R.w4[2], v);
        v = __dp4a((int)yv.w, (int)R.w4[3], v);
        int m = max(v, R.thr4);
        m += __shfl_xor_sync(0xffffffffu, m, 1);
        m += __shfl_xor_sync(0xffffffffu, m, 2);
        m += __shfl_xor_sync(0xffffffffu, m, 4);
        if (e8 == (q & 7) && k < F2) {
            const int o = eclip8(et_div(m + R.off4, R.fac4, R.m4, R.sh4, R.sg4, mg));
            y4[k * T64 + q] = (int8_t)o;
            if (tap4) tap4[((size_t)b * F2 + k) * T64 + q] = (int8_t)o;
        }
    }
    et_bar_sync_l();
    for (int n = tid >> 5; n < N; n += kEtL) {
        const int8_t* w = net->w5 + (size_t)n * F2 * T64;
        int z = 0;
        for (int i = tid & 31; i < F2 * T64; i += 32) z += (int)y4[i] * (int)w[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
        if ((tid & 31) == 0) logits[(size_t)b * N + n] = (int8_t)eclip8(et_div(z + net->b5[n], net->fac5, net->mg5, net->sh5, net->sg5, mg));
    }
    et_bar_sync_l();
}

__global__ void __launch_bounds__(kEtThreads, 1)
eegnet_tc_kernel(const EegNet* __restrict__ net, const int8_t* __restrict__ x, long long B, int8_t* __restrict__ logits,
                 int8_t* __restrict__ tap2, int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(1024) uint8_t esm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const uint32_t sbase = smem_u32(esm);
    const int C = net->C, T = net->T, K2 = net->K2, T8 = net->T8;
    const int NCH = net->NCH, P = net->P, CT = net->CT, NXS = net->NXS, KC = 16 * CT;
    const int y2s = et_y2s(T8), PL = 64;
    const EtSmem sm = et_smem(C, T, K2, net->F2, NCH, P, CT, NXS);
    const int xs_bytes = 16 * CT * P, a2_bytes = 8 * KC * 128;
    const long long first = blockIdx.x, stride = gridDim.x;
    const int ntr = first < B ? (int)((B - first + stride - 1) / stride) : 0;

    // ---- prologue
    if (tid == 0) {
        for (int i = 0; i < 2; i++) {
            mbar_init(sbase + kEtXsFull + 8 * i, kEtS);
            mbar_init(sbase + kEtXsEmpty + 8 * i, 1);            // commit of M
            mbar_init(sbase + kEtA2Full + 8 * i, 4 * CT);        // the four warps of every tile of the chunk
            mbar_init(sbase + kEtA2Empty + 8 * i, 1);            // commit of M
            mbar_init(sbase + kEtT2Full + 8 * i, 1);             // commit of M
            mbar_init(sbase + kEtT2Empty + 8 * i, kEtE2);
            mbar_init(sbase + kEtY2Full + 8 * i, kEtE2);
            mbar_init(sbase + kEtY2Empty + 8 * i, 1);
        }
        for (int i = 0; i < kEtSets; i++) {
            mbar_init(sbase + kEtT1Full + 8 * i, 1);             // commit of M
            mbar_init(sbase + kEtT1Empty + 8 * i, 4);
        }
        mbar_fence_init();
    }
    for (int i = tid; i < kEtB1Bytes / 16; i += kEtThreads)
        reinterpret_cast<uint4*>(esm + kEtOffB1)[i] = reinterpret_cast<const uint4*>(net->b1umma)[i];
    for (int i = tid; i < 16 * 8 * KC / 16; i += kEtThreads)
        reinterpret_cast<uint4*>(esm + sm.off_b2)[i] = reinterpret_cast<const uint4*>(net->b2umma)[i];
    {   // constants: [0..7] m32, [8..23] offm32 (lo, hi), [24..39] thr2, [40..55] off2, [56..71] fac2
        int* cw = reinterpret_cast<int*>(esm + kEtOffCons);
        for (int i = tid; i < 120; i += kEtThreads) {     // ... [72..87] mg2, [88..103] sh2, [104..119] sg2
            int v = 0;
            if (i < 8) v = net->m32[i];
            else if (i < 24) v = (int)((unsigned long long)net->offm32[(i - 8) >> 1] >> (32 * ((i - 8) & 1)));
            else if (i < 40) v = (i - 24) < K2 ? -(net->off2[i - 24] >> 3) : 0;
            else if (i < 56) v = (i - 40) < K2 ? net->off2[i - 40] : 0;
            else if (i < 72) v = (i - 56) < K2 ? net->fac2[i - 56] : 1;
            else if (i < 88) v = (i - 72) < K2 ? net->mg2[i - 72] : 0;
            else if (i < 104) v = (i - 88) < K2 ? net->sh2[i - 88] : 0;
            else v = (i - 104) < K2 ? net->sg2[i - 104] : 1;
            cw[i] = v;
        }
    }
    {
        uint4* z = reinterpret_cast<uint4*>(esm + sm.off_a2);     // A2 rows of channels >= C, the zero padding of the input rows and of y2
        for (int i = tid; i < (sm.total - sm.off_a2) / 16; i += kEtThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (warp == kEtWarpM) tmem_alloc(sbase + kEtTmemSlot, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<const uint32_t*>(esm + kEtTmemSlot);
    const int* cw = reinterpret_cast<const int*>(esm + kEtOffCons);
    EPROF_BEGIN

    if (warp < kEtL) {
        // ===================================================================================== L: layers 3 - 5
        EtLRegs R;
        et_l_setup(net, tid, R);
        for (int i = 0; i < ntr; i++) {
            const long long b = first + (long long)i * stride;
            const int yb = i & 1;
            EWAITS(0, sbase + kEtY2Full + 8 * yb, (i >> 1) & 1);
            et_layers345(net, R, reinterpret_cast<const int8_t*>(esm + sm.off_y2 + yb * K2 * y2s), y2s, reinterpret_cast<int8_t*>(esm + sm.off_y3),
                         reinterpret_cast<int8_t*>(esm + sm.off_y4), b, logits, tap3, tap4, tid);
            if (tid == 0) mbar_arrive(sbase + kEtY2Empty + 8 * yb);
        }
    } else if (warp < kEtWarpE1) {
        // ===================================================================================== E2: batch_norm2 + ReLU + pool + quant3
        const int qi = warp & 3;
        int thr[4];
#pragma unroll
        for (int j = 0; j < 4; j++) thr[j] = cw[24 + 8 * (j >> 1) + 2 * q + (j & 1)];      // column 8 nt + 2 q + e, j = 2 nt + e
        // the reduction over the 8 rows g of a pooling group halves the values a lane carries in every round (xor 4, 8, 16 on the lane
        // = bits 0, 1, 2 of g): lane g ends with the two columns e = 0, 1 of nt = g & 1 in the group rg = 2 ((g >> 1) & 1) + ((g >> 2) & 1)
        const int nt_l = g & 1, rg_l = 2 * ((g >> 1) & 1) + ((g >> 2) & 1);
        const int k0 = 8 * nt_l + 2 * q;
        const int off_a = cw[40 + k0], off_b = cw[41 + k0], fac_a = cw[56 + k0], fac_b = cw[57 + k0];
        const int mg_a = cw[72 + k0], mg_b = cw[73 + k0], sh_a = cw[88 + k0], sh_b = cw[89 + k0], sg_a = cw[104 + k0], sg_b = cw[105 + k0];
        const int mgc = net->magic_rest;
        for (int i = 0; i < ntr; i++) {
            const long long b = first + (long long)i * stride;
            const int yb = i & 1;
            EWAITS(1, sbase + kEtY2Empty + 8 * yb, ((i >> 1) & 1) ^ 1);       // the L warps are done with this y2 buffer (trial i - 2)
            int8_t* y2 = reinterpret_cast<int8_t*>(esm + sm.off_y2 + yb * K2 * y2s);
            for (int ch = 0; ch < NCH; ch++) {
                const int gch = i * NCH + ch, buf = gch & 1;
                EWAITS(0, sbase + kEtT2Full + 8 * buf, (gch >> 1) & 1);
                tc_fence_after();
                uint32_t r0[8], r1[8];
                const uint32_t ta = tmem_base + ((uint32_t)(32 * qi) << 16) + kEtTmT2 + 16 * buf;
                tmem_ld_16x256b_x2(r0, ta);
                tmem_ld_16x256b_x2(r1, ta + (16u << 16));
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(sbase + kEtT2Empty + 8 * buf);
                // r[4 nt + c] = (row g + 8 (c >> 1), column 8 nt + 2 q + (c & 1)): threshold, then sum over the 8 rows g of a group
                // v[8 nt + 4 b2 + 2 c2 + e]: row group rg = 2 b2 + c2, column 8 nt + 2 q + e
                int v[16];
#pragma unroll
                for (int nt = 0; nt < 2; nt++)
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const int e = c & 1, th = thr[2 * nt + e];
                        v[8 * nt + 2 * (c >> 1) + e] = max((int)r0[4 * nt + c], th);           // rg = c >> 1      (b2 = 0)
                        v[8 * nt + 4 + 2 * (c >> 1) + e] = max((int)r1[4 * nt + c], th);       // rg = 2 + (c >> 1) (b2 = 1)
                    }
                {
                    const bool up = g & 1;                  // round 1 (lane xor 4): keep nt = g & 1
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const int send = up ? v[j] : v[j + 8], keep = up ? v[j + 8] : v[j];
                        v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                    }
                }
                {
                    const bool up = (g >> 1) & 1;           // round 2 (lane xor 8): keep b2
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int send = up ? v[j] : v[j + 4], keep = up ? v[j + 4] : v[j];
                        v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                    }
                }
                {
                    const bool up = (g >> 2) & 1;           // round 3 (lane xor 16): keep c2
#pragma unroll
                    for (int j = 0; j < 2; j++) {
                        const int send = up ? v[j] : v[j + 2], keep = up ? v[j + 2] : v[j];
                        v[j] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                    }
                }
                const int sa = v[0], sb = v[1];
                const int pg = 16 * ch + 4 * qi + rg_l;                          // pooled sample of the trial
                if (pg < T8) {
                    if (k0 < K2) {
                        const int o = eclip8(et_div(sa + off_a, fac_a, mg_a, sh_a, sg_a, mgc));
                        y2[k0 * y2s + 7 + pg] = (int8_t)o;
                        if (tap2) tap2[((size_t)b * K2 + k0) * T8 + pg] = (int8_t)o;
                    }
                    if (k0 + 1 < K2) {
                        const int o = eclip8(et_div(sb + off_b, fac_b, mg_b, sh_b, sg_b, mgc));
                        y2[(k0 + 1) * y2s + 7 + pg] = (int8_t)o;
                        if (tap2) tap2[((size_t)b * K2 + k0 + 1) * T8 + pg] = (int8_t)o;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + kEtY2Full + 8 * yb);
        }
    } else if (warp < kEtWarpS) {
        // ===================================================================================== E1: batch_norm1 / quant2 -> A operand of MMA 2
        const int es = (warp - kEtWarpE1) >> 2, qi = warp & 3;
        const int r = lane & 7, cl = 4 * qi + (lane >> 3);
        int m1c[8];
        long long offm1c[8];
#pragma unroll
        for (int f = 0; f < 8; f++) {
            m1c[f] = cw[f];
            offm1c[f] = *reinterpret_cast<const long long*>(cw + 8 + 2 * f);
        }
        // this set owns the tiles es, es + kEtSets, ... of the CTA's tile sequence (trial, chunk, channel tile); no division per tile
        int i = 0, ch = 0, ct = es;
        while (ct >= CT) { ct -= CT; ch++; }
        while (ch >= NCH) { ch -= NCH; i++; }
        for (int use = 0; i < ntr; use++) {
                {
                    const int gch = i * NCH + ch, buf = gch & 1;
                    EWAIT(0, sbase + kEtT1Full + 8 * es, use & 1);
                    EWAIT(1, sbase + kEtA2Empty + 8 * buf, ((gch >> 1) & 1) ^ 1);     // MMA 2 of chunk gch - 2 has read this buffer
                    tc_fence_after();
                    // a last tile of at most 8 channels is an M = 64 instruction (MIBMI_ET_M64): its row i sits in TMEM lane
                    // 32 (i / 16) + i % 16, i.e. a lane quarter holds two channels in its lanes 0 .. 15
                    const bool half = MIBMI_ET_M64 && C - 16 * ct <= 8;
                    const int c = half ? 16 * ct + 2 * qi + ((lane >> 3) & 1) + (lane >= 16 ? 64 : 0) : 16 * ct + cl;
                    if (half ? (16 * ct + 2 * qi < C) : (16 * ct + 4 * qi < C)) {       // warp-uniform: some channel of the quarter exists
                        const uint32_t ta = tmem_base + ((uint32_t)(32 * qi) << 16) + 128 * es;
                        uint8_t* a2 = esm + sm.off_a2 + buf * a2_bytes + c * 128 + ((r ^ (c & 7)) << 4);
#pragma unroll
                        for (int fp = 0; fp < 4; fp++) {
                            uint32_t ra[16], rb[16];
                            et_tmem_ld_32x32b_x16(ra, ta + 32 * fp);
                            et_tmem_ld_32x32b_x16(rb, ta + 32 * fp + 16);
                            tmem_ld_wait();
#pragma unroll
                            for (int h = 0; h < 2; h++) {
                                const int f = 2 * fp + h;
                                const int m = m1c[f];
                                const long long offm = offm1c[f];
                                uint32_t w[4];
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    int v[4];
#pragma unroll
                                    for (int e = 0; e < 4; e++) v[e] = et_div_fold((int)(h ? rb[4 * j + e] : ra[4 * j + e]), m, offm);
                                    w[j] = pack4_sat(v[0], v[1], v[2], v[3]);
                                }
                                if (c < C) *reinterpret_cast<uint4*>(a2 + f * KC * 128) = make_uint4(w[0], w[1], w[2], w[3]);
                            }
                        }
                    }
                    tc_fence_before();
                    fence_proxy_async();          // A2 is read by the tensor core (async proxy)
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive(sbase + kEtT1Empty + 8 * es);
                        mbar_arrive(sbase + kEtA2Full + 8 * buf);
                    }
                }
                ct += kEtSets;
                while (ct >= CT) { ct -= CT; ch++; }
                while (ch >= NCH) { ch -= NCH; i++; }
        }
    } else if (warp < kEtWarpM) {
        // ===================================================================================== S: trial -> padded rows (x[c][t] at byte PL + t)
        // PL = 64: every row starts on a 16-byte boundary of the stage.  Rows that are 16-byte aligned in global memory too (64 x 480)
        // arrive as one bulk copy each (TMA, no instructions per byte); otherwise a lane builds a 16-byte chunk from five aligned words
        // (funnel shift) and stores it once; bytes >= T of the last chunk are written as zeros (they are the convolution's padding)
        const int sw = warp - kEtWarpS;
        const int chunks = (T + 15) >> 4;
        const uintptr_t gend = (reinterpret_cast<uintptr_t>(x) + (size_t)B * C * T + 3) & ~(uintptr_t)3;
        for (int i = 0; i < ntr; i++) {
            const long long b = first + (long long)i * stride;
            const int st = NXS == 1 ? 0 : (i & 1), xuse = NXS == 1 ? i : (i >> 1);       // NXS is 1 or 2
            EWAITS(0, sbase + kEtXsEmpty + 8 * st, (xuse & 1) ^ 1);
            uint8_t* xs = esm + sm.off_xs + st * xs_bytes + PL;
            const int8_t* xb = x + (size_t)b * C * T;
            const uint32_t bar = sbase + kEtXsFull + 8 * st;
            if (((reinterpret_cast<uintptr_t>(xb) | (uintptr_t)T) & 15) == 0) {
                if (sw == 0) {
                    if (lane == 0) mbar_expect_tx(bar, (uint32_t)(C * T));
                    __syncwarp();
                    for (int c = lane; c < C; c += 32) bulk_g2s(smem_u32(xs + c * P), xb + (size_t)c * T, (uint32_t)T, bar);
                } else if (lane == 0) {
                    mbar_arrive(bar);
                }
            } else {
                for (int c = sw; c < C; c += kEtS) {
                    const uintptr_t ra = reinterpret_cast<uintptr_t>(xb) + (size_t)c * T;
                    const int sh = 8 * (int)(ra & 3);
                    uint8_t* drow = xs + c * P;
                    for (int j0 = 0; j0 < chunks; j0 += 96) {
                        uint32_t w[3][5];
#pragma unroll
                        for (int pp = 0; pp < 3; pp++) {
                            const int j = j0 + 32 * pp + lane;
                            const uintptr_t wa = (ra & ~(uintptr_t)3) + 16 * (size_t)j;
#pragma unroll
                            for (int u = 0; u < 5; u++)
                                w[pp][u] = (j < chunks && wa + 4 * u < gend) ? __ldg(reinterpret_cast<const uint32_t*>(wa + 4 * u)) : 0u;
                        }
#pragma unroll
                        for (int pp = 0; pp < 3; pp++) {
                            const int j = j0 + 32 * pp + lane;
                            if (j < chunks) {
                                uint32_t o[4];
#pragma unroll
                                for (int u = 0; u < 4; u++) {
                                    o[u] = __funnelshift_r(w[pp][u], w[pp][u + 1], sh);
                                    const int nv = T - 16 * j - 4 * u;                 // valid bytes of this word
                                    if (nv < 4) o[u] = nv <= 0 ? 0u : (o[u] & (0xffffffffu >> (8 * (4 - nv))));
                                }
                                *reinterpret_cast<uint4*>(drow + 16 * j) = make_uint4(o[0], o[1], o[2], o[3]);
                            }
                        }
                    }
                }
                fence_proxy_async();              // the rows are read by the tensor core (async proxy)
                __syncwarp();
                if (lane == 0) mbar_arrive(bar);
            }
        }
    } else {
        // ===================================================================================== M: both MMAs
        constexpr uint32_t idesc1 = et_idesc_i8_kk(128, 128), idesc1h = et_idesc_i8_kk(64, 128), idesc2 = umma_idesc_i8(128, 16);
        const uint64_t da1 = umma_desc(sbase + sm.off_xs, 16, (uint32_t)P, 0), db1 = umma_desc(sbase + kEtOffB1, 128, 1280, 0);
        const uint64_t da2 = umma_desc(sbase + sm.off_a2, 16, 1024, 2), db2 = umma_desc(sbase + sm.off_b2, 128, (uint32_t)(8 * KC / 16) * 128, 0);
        const uint32_t a1_hi = (uint32_t)(da1 >> 32), b1_hi = (uint32_t)(db1 >> 32), a2_hi = (uint32_t)(da2 >> 32), b2_hi = (uint32_t)(db2 >> 32);
        const int ks2 = 8 * KC / 32;
        auto mma2 = [&](int gch) {                // spatial conv of chunk gch (issued one chunk late: E1 needs the time)
            const int buf = gch & 1;
            EWAIT(1, sbase + kEtA2Full + 8 * buf, (gch >> 1) & 1);
            EWAIT(2, sbase + kEtT2Empty + 8 * buf, ((gch >> 1) & 1) ^ 1);
            tc_fence_after();
            if (elect_one()) {
                for (int ks = 0; ks < ks2; ks++)
                    umma_i8_lohi(tmem_base + kEtTmT2 + 16 * buf, (uint32_t)da2 + ((buf * a2_bytes + ks * 4096) >> 4), a2_hi,
                                 (uint32_t)db2 + ((ks * 256) >> 4), b2_hi, idesc2, ks > 0);
                umma_commit(sbase + kEtT2Full + 8 * buf);
                umma_commit(sbase + kEtA2Empty + 8 * buf);
            }
            __syncwarp();
        };
        int tb = 0, use = 0, st = 0, xuse = 0;             // accumulator buffer of the next tile / its use count; input stage / its use count
        for (int i = 0; i < ntr; i++) {
            EWAIT(0, sbase + kEtXsFull + 8 * st, xuse & 1);
            for (int ch = 0; ch < NCH; ch++) {
                const int gch = i * NCH + ch;
                for (int ct = 0; ct < CT; ct++) {
                    EWAIT(2, sbase + kEtT1Empty + 8 * tb, (use & 1) ^ 1);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t a_lo = (uint32_t)da1 + ((st * xs_bytes + 16 * ct * P + 128 * ch) >> 4);
#pragma unroll
                        for (int s = 0; s < 5; s++)
                            umma_i8_lohi(tmem_base + 128 * tb, a_lo + ((32 * s) >> 4), a1_hi, (uint32_t)db1 + ((256 * s) >> 4), b1_hi,
                                         (MIBMI_ET_M64 && C - 16 * ct <= 8) ? idesc1h : idesc1, s > 0);
                        umma_commit(sbase + kEtT1Full + 8 * tb);
                        if (ch == NCH - 1 && ct == CT - 1) umma_commit(sbase + kEtXsEmpty + 8 * st);
                    }
                    __syncwarp();
                    if (++tb == kEtSets) { tb = 0; use++; }
                }
                if (gch > 0) mma2(gch - 1);
            }
            if (++st == NXS) { st = 0; xuse++; }
        }
        if (ntr > 0) mma2(ntr * NCH - 1);
    }
    EPROF_END
    tc_fence_before();
    __syncthreads();
    if (warp == kEtWarpM) tmem_dealloc(tmem_base, 512);
}

}  // namespace mibmi

using namespace mibmi;

struct mibmi_eegnet {
    int device = 0, num_sms = 0;
    int C = 0, T = 0, K2 = 0, F2 = 0, N = 0, D = 0, threads = 256, magic1 = 0;
    int F1 = 0, KS = 0;                 // tensor-core kernel: warps = F1, k-steps; KS == 0: no instantiation (dp4a kernel)
    size_t smem = 0, smem_mma = 0, smem_tc = 0;
    int tc = 0;                         // tcgen05 kernel (eegnet_tc_kernel) available for this network
    EegNet* net_dev = nullptr;
    int8_t* w5_dev = nullptr;
    uint32_t* w1frag_dev = nullptr;
    uint8_t* b1_dev = nullptr;          // operands of the tcgen05 kernel
    uint8_t* b2_dev = nullptr;
    std::atomic<int> last_kernel{0};    // 0 none, 1 eegnet_tc_kernel, 2 eegnet_mma_kernel, 3 eegnet_forward_kernel
    std::atomic<long long> launches{0};
    std::atomic<int> last_cuda{0};
};

namespace {
struct Guard {
    int prev = -1;
    cudaError_t err;
    explicit Guard(int dev) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
    }
    ~Guard() {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
};
}  // namespace

extern "C" {

int mibmi_eegnet_create(const mibmi_eegnet_params_t* p, int device, mibmi_eegnet_t** out) {
    if (!p || !out) return MIBMI_ERR_NULL;
    *out = nullptr;
    if (!p->l1_weight || !p->l1_factor || !p->l1_offset || !p->l2_weight || !p->l2_factor || !p->l2_offset || !p->l3_weight ||
        !p->l4_weight || !p->l4_factor || !p->l4_offset || !p->l5_weight || !p->l5_bias)
        return MIBMI_ERR_NULL;
    const int C = p->C, T = p->T, F1 = p->F1, D = p->D, K2 = p->F1 * p->D, F2 = p->F2, N = p->N, Nf = p->Nf;
    if (C < 1 || C > kMaxC || T < 64 || F1 < 4 || F1 > kEegMaxF1 || F1 % 4 || (D != 1 && D != 2) || F2 < 1 || F2 > kMaxF2 ||
        N < 1 || N > kMaxN || Nf < 4 || Nf > kEegMaxNf || Nf % 4)
        return MIBMI_ERR_SHAPE;
    EegNet* h = new (std::nothrow) EegNet;
    if (!h) return MIBMI_ERR_CUDA;
    memset(h, 0, sizeof(EegNet));
    h->C = C; h->T = T; h->F1 = F1; h->D = D; h->K2 = K2; h->F2 = F2; h->N = N; h->Nf = Nf;
    h->T8 = T / 8; h->T64 = h->T8 / 8;
    h->XP = (T + Nf + 8 + 3) / 4 * 4;
    h->G = (T + 3) / 4;
    h->NW = Nf / 4 + 1;
    const int items = (F1 / 4) * h->G, rounds = (items + 383) / 384;      // layer-1/2 work items per trial -> whole rounds
    const int threads = ((items + rounds - 1) / rounds + 31) / 32 * 32;
    bool zero = p->l3_factor == 0 || p->l5_factor == 0;
    for (int f = 0; f < F1; f++) {
        h->fac1[f] = p->l1_factor[f]; h->off1[f] = p->l1_offset[f];
        zero |= h->fac1[f] == 0;
        for (int i = 0; i < h->NW; i++)
            for (int s = 0; s < 4; s++) {
                uint32_t w = 0;
                for (int e = 0; e < 4; e++) {
                    const int j = 4 * i + e - s;
                    const int v = (j >= 0 && j < Nf) ? p->l1_weight[f * Nf + j] : 0;
                    w |= (uint32_t)(uint8_t)(int8_t)v << (8 * e);
                }
                h->w1s[(f * h->NW + i) * 4 + s] = (int)w;
            }
    }
    h->magic1 = 1;
    for (int f = 0; f < F1; f++) {
        long long sum = 0;
        for (int j = 0; j < Nf; j++) sum += p->l1_weight[f * Nf + j] < 0 ? -(long long)p->l1_weight[f * Nf + j] : p->l1_weight[f * Nf + j];
        const long long fa = h->fac1[f] < 0 ? -(long long)h->fac1[f] : h->fac1[f];
        const long long oa = h->off1[f] < 0 ? -(long long)h->off1[f] : h->off1[f];
        h->sg1[f] = h->fac1[f] < 0 ? -1 : 1;
        // tdiv(x, d) == (mulhi(x, m) >> s) + (x < 0) for d >= 4, |x| < 2^29; otherwise hardware division
        bool has_m128 = false;
        for (int j = 0; j < Nf; j++) has_m128 |= p->l1_weight[f * Nf + j] == -128;       // cannot be negated in int8
        if (fa < 4 || fa > 0x7fffffffLL || 128 * sum + oa >= (1LL << 29) || (has_m128 && h->fac1[f] < 0)) { h->magic1 = 0; continue; }
        int lg = 0;
        while ((1LL << (lg + 1)) <= fa) lg++;
        const long long mm = (long long)(((__int128)1 << (30 + lg)) / fa) + 1;
        if (mm >= (1LL << 31)) { h->magic1 = 0; continue; }
        h->mg1[f] = (int)mm;
        h->sh1[f] = lg - 2;
    }
    for (int k = 0; k < K2; k++) {
        h->fac2[k] = p->l2_factor[k]; h->off2[k] = p->l2_offset[k];
        zero |= h->fac2[k] == 0;
    }
    for (int k = 0; k < F2; k++) {
        h->fac4[k] = p->l4_factor[k]; h->off4[k] = p->l4_offset[k];
        zero |= h->fac4[k] == 0;
    }
    if (zero) { delete h; return MIBMI_ERR_PARAM; }
    h->fac3 = p->l3_factor; h->fac5 = p->l5_factor;
    for (int n = 0; n < N; n++) h->b5[n] = p->l5_bias[n];
    memcpy(h->w2, p->l2_weight, (size_t)K2 * C);
    memcpy(h->w3, p->l3_weight, (size_t)K2 * kL3Taps);
    memcpy(h->w4, p->l4_weight, (size_t)F2 * K2);

    mibmi_eegnet* n = new (std::nothrow) mibmi_eegnet;
    if (!n) { delete h; return MIBMI_ERR_CUDA; }
    n->device = device; n->C = C; n->T = T; n->K2 = K2; n->F2 = F2; n->N = N; n->D = D;
    n->threads = threads < 128 ? 128 : threads;
    n->magic1 = h->magic1;
    n->smem = (size_t)F1 * h->NW * 16 + (size_t)C * h->XP + (size_t)K2 * h->G * 4 + (size_t)K2 * (h->T8 + 15) + (size_t)K2 * h->T8 +
              (size_t)F2 * h->T64 + 16;
    Guard dg(device);
    cudaError_t e = dg.err;
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess && n->smem > (size_t)prop.sharedMemPerBlockOptin) { delete h; delete n; return MIBMI_ERR_SHAPE; }
    if (e == cudaSuccess) n->num_sms = prop.multiProcessorCount;
    const size_t w5n = (size_t)N * F2 * h->T64;
    if (e == cudaSuccess) e = cudaMalloc(&n->w5_dev, w5n ? w5n : 1);
    if (e == cudaSuccess && w5n) e = cudaMemcpy(n->w5_dev, p->l5_weight, w5n, cudaMemcpyHostToDevice);
    h->w5 = n->w5_dev;
    // tensor-core kernel: Toeplitz B fragments of every filter (see eegnet_mma_kernel)
    n->F1 = F1;
    h->KS = (Nf + 63) / 32;
    h->MT = ((T + 31) / 32 + 15) / 16;
    {
        const int need1 = 32 * 16 * h->MT + 32 * h->KS, need2 = T + Nf + 32;
        h->XPM = ((need1 > need2 ? need1 : need2) + 15) / 16 * 16;
    }
    const bool mma_shape = Nf % 32 == 0 && (h->KS == 2 || h->KS == 3 || h->KS == 5);
    n->smem_mma = (size_t)C * h->XPM + (size_t)K2 * (h->T8 + 15) + (size_t)K2 * h->T8 + (size_t)F2 * h->T64 + 16;
    if (e == cudaSuccess && mma_shape && n->smem_mma <= (size_t)prop.sharedMemPerBlockOptin) {
        const size_t nw = (size_t)F1 * 4 * h->KS * 2 * 32;
        uint32_t* fr = new (std::nothrow) uint32_t[nw];
        if (fr) {
            for (int f = 0; f < F1; f++)
                for (int v = 0; v < 4; v++)
                    for (int s = 0; s < h->KS; s++)
                        for (int hh = 0; hh < 2; hh++)
                            for (int lane = 0; lane < 32; lane++) {
                                const int g = lane >> 2, q = lane & 3;
                                uint32_t w = 0;
                                for (int i = 0; i < 4; i++) {
                                    const int tap = 32 * s + 16 * hh + 4 * q + i - (8 * v + g);
                                    const int sg = h->magic1 ? h->sg1[f] : 1;       // tdiv(x, -d) == tdiv(-x, d)
                                    const int val = (tap >= 0 && tap < Nf) ? sg * p->l1_weight[f * Nf + tap] : 0;
                                    w |= (uint32_t)(uint8_t)(int8_t)val << (8 * i);
                                }
                                fr[(((f * 4 + v) * h->KS + s) * 2 + hh) * 32 + lane] = w;
                            }
            e = cudaMalloc(&n->w1frag_dev, nw * 4);
            if (e == cudaSuccess) e = cudaMemcpy(n->w1frag_dev, fr, nw * 4, cudaMemcpyHostToDevice);
            delete[] fr;
            if (e == cudaSuccess) n->KS = h->KS;
        }
    }
    h->w1frag = n->w1frag_dev;
    // tcgen05 kernel: F1 = 8 filters of 128 taps, every layer-1 factor inside the range of the "no shift" magic
    if (e == cudaSuccess && F1 == 8 && Nf == 128 && C <= 64 && K2 <= 16 && F2 <= 16 && h->magic1) {
        bool ok = true;
        for (int f = 0; f < F1 && ok; f++) {
            long long sum = 0;
            for (int j = 0; j < Nf; j++) sum += p->l1_weight[f * Nf + j] < 0 ? -(long long)p->l1_weight[f * Nf + j] : p->l1_weight[f * Nf + j];
            const long long d = h->fac1[f] < 0 ? -(long long)h->fac1[f] : h->fac1[f];
            const long long oa = h->off1[f] < 0 ? -(long long)h->off1[f] : h->off1[f];
            const long long xmax = 128 * sum + oa;
            const long long mm = (1LL << 32) / d + 1, er = mm * d - (1LL << 32), lim = xmax < 130 * d ? xmax : 130 * d;
            if (d < 4 || mm >= (1LL << 31) || (__int128)lim * er >= ((__int128)1 << 32)) { ok = false; break; }
            h->m32[f] = (int)mm;
            h->offm32[f] = (long long)h->sg1[f] * h->off1[f] * mm;
        }
        {   // layers 2 - 5: tdiv(x, d) == (mulhi(x, m) >> s) + (x < 0) for |d| >= 4, |x| < 2^29, sign of d applied after
            auto magic = [](long long fac, long long xmax, int* m, int* sh, int* sg) -> bool {
                const long long fa = fac < 0 ? -fac : fac;
                *sg = fac < 0 ? -1 : 1;
                if (fa < 4 || fa > 0x7fffffffLL || xmax >= (1LL << 29)) return false;
                int lg = 0;
                while ((1LL << (lg + 1)) <= fa) lg++;
                const long long mm = (long long)(((__int128)1 << (30 + lg)) / fa) + 1;
                if (mm >= (1LL << 31)) return false;
                *m = (int)mm;
                *sh = lg - 2;
                return true;
            };
            auto l1 = [](const int8_t* w, int n) { long long t = 0; for (int i = 0; i < n; i++) t += w[i] < 0 ? -(long long)w[i] : w[i]; return t; };
            auto ab = [](long long v) { return v < 0 ? -v : v; };
            bool mr = true;
            for (int k = 0; k < K2 && mr; k++) {
                const long long am = 128 * l1(p->l2_weight + k * C, C), th = ab(-((long long)h->off2[k] >> 3));
                mr = magic(h->fac2[k], 8 * (am > th ? am : th) + ab(h->off2[k]), &h->mg2[k], &h->sh2[k], &h->sg2[k]);
            }
            long long m3x = 0;
            for (int k = 0; k < K2; k++) { const long long v = 128 * l1(p->l3_weight + k * kL3Taps, kL3Taps); m3x = v > m3x ? v : m3x; }
            mr = mr && magic(h->fac3, m3x, &h->mg3, &h->sh3, &h->sg3);
            for (int k = 0; k < F2 && mr; k++) {
                const long long am = 128 * l1(p->l4_weight + k * K2, K2), th = ab(-((long long)h->off4[k] >> 3));
                mr = magic(h->fac4[k], 8 * (am > th ? am : th) + ab(h->off4[k]), &h->mg4[k], &h->sh4[k], &h->sg4[k]);
            }
            long long m5x = 0;
            for (int nn = 0; nn < N; nn++) {
                const long long v = 128 * l1(p->l5_weight + (size_t)nn * F2 * h->T64, F2 * h->T64) + ab(h->b5[nn]);
                m5x = v > m5x ? v : m5x;
            }
            mr = mr && magic(h->fac5, m5x, &h->mg5, &h->sh5, &h->sg5);
            h->magic_rest = mr ? 1 : 0;
        }
        h->NCH = (T + 127) / 128;
        h->P = 128 * h->NCH + 144;
        h->CT = (C + 15) / 16;
        h->NXS = 2;
        if (ok && (size_t)et_smem(C, T, K2, F2, h->NCH, h->P, h->CT, 2).total > (size_t)prop.sharedMemPerBlockOptin) h->NXS = 1;
        if (ok && (size_t)et_smem(C, T, K2, F2, h->NCH, h->P, h->CT, h->NXS).total > (size_t)prop.sharedMemPerBlockOptin) ok = false;
        if (ok) {
            const int KC = 16 * h->CT, K2T = 8 * KC;
            const size_t nb1 = kEtB1Bytes, nb2 = (size_t)16 * K2T;
            uint8_t* b1 = new (std::nothrow) uint8_t[nb1];
            uint8_t* b2 = new (std::nothrow) uint8_t[nb2];
            if (b1 && b2) {
                memset(b1, 0, nb1);
                memset(b2, 0, nb2);
                for (int nn = 0; nn < 128; nn++)            // B1[(f, n)][kk] = sg w1[f][kk - n - 1]: output n of a row reads bytes n + 1 .. n + 128 of its window
                    for (int kk = 0; kk < 160; kk++) {
                        const int f = nn >> 4, tap = kk - (nn & 15) - 1;     // x[t] sits at byte 64 + t of its padded row, the window of output t starts at byte t
                        if (tap >= 0 && tap < Nf)
                            b1[(nn >> 3) * 1280 + (kk >> 4) * 128 + (nn & 7) * 16 + (kk & 15)] = (uint8_t)(int8_t)(h->sg1[f] * p->l1_weight[f * Nf + tap]);
                    }
                for (int k = 0; k < K2; k++)                // B2[k][(f, c)] = w2[k][c] for f = k / D (depthwise: groups = F1)
                    for (int c = 0; c < C; c++) {
                        const int kk = (k / D) * KC + c;
                        b2[(k >> 3) * (K2T / 16) * 128 + (kk >> 4) * 128 + (k & 7) * 16 + (kk & 15)] = (uint8_t)p->l2_weight[k * C + c];
                    }
                e = cudaMalloc(&n->b1_dev, nb1);
                if (e == cudaSuccess) e = cudaMalloc(&n->b2_dev, nb2);
                if (e == cudaSuccess) e = cudaMemcpy(n->b1_dev, b1, nb1, cudaMemcpyHostToDevice);
                if (e == cudaSuccess) e = cudaMemcpy(n->b2_dev, b2, nb2, cudaMemcpyHostToDevice);
                if (e == cudaSuccess) {
                    n->tc = 1;
                    n->smem_tc = (size_t)et_smem(C, T, K2, F2, h->NCH, h->P, h->CT, h->NXS).total;
                }
            }
            delete[] b1;
            delete[] b2;
        }
    }
    h->b1umma = n->b1_dev;
    h->b2umma = n->b2_dev;
    if (e == cudaSuccess) e = cudaMalloc(&n->net_dev, sizeof(EegNet));
    if (e == cudaSuccess) e = cudaMemcpy(n->net_dev, h, sizeof(EegNet), cudaMemcpyHostToDevice);
    delete h;
    if (e != cudaSuccess) {
        if (n->w5_dev) cudaFree(n->w5_dev);
        if (n->w1frag_dev) cudaFree(n->w1frag_dev);
        if (n->b1_dev) cudaFree(n->b1_dev);
        if (n->b2_dev) cudaFree(n->b2_dev);
        if (n->net_dev) cudaFree(n->net_dev);
        delete n;
        return MIBMI_ERR_CUDA;
    }
    *out = n;
    return MIBMI_OK;
}

int mibmi_eegnet_destroy(mibmi_eegnet_t* n) {
    if (!n) return MIBMI_OK;
    {
        Guard dg(n->device);
        if (n->w5_dev) cudaFree(n->w5_dev);
        if (n->w1frag_dev) cudaFree(n->w1frag_dev);
        if (n->b1_dev) cudaFree(n->b1_dev);
        if (n->b2_dev) cudaFree(n->b2_dev);
        if (n->net_dev) cudaFree(n->net_dev);
    }
    delete n;
    return MIBMI_OK;
}

int mibmi_eegnet_forward_int8_taps(mibmi_eegnet_t* n, const int8_t* x_dev, int64_t B, int8_t* logits_dev, int8_t* y2, int8_t* y3,
                                   int8_t* y4, void* stream) {
    if (!n) return MIBMI_ERR_NULL;
    if (B < 0) return MIBMI_ERR_PARAM;
    if (B == 0) return MIBMI_OK;
    if (!x_dev || !logits_dev) return MIBMI_ERR_NULL;
    Guard dg(n->device);
    cudaError_t e = dg.err;
    const char* force = getenv("MIBMI_EEGNET_DP4A");                       // experiment switch: CUDA-core kernel
    const char* notc = getenv("MIBMI_EEGNET_TC");                          // experiment switch: "0" = the mma.sync kernel
    int which = 0;
    if (n->tc && !(notc && notc[0] == '0') && !(force && force[0] == '1')) {
        auto kern = eegnet_tc_kernel;        // D only shapes the block-diagonal operand of the second MMA (b2umma)
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)n->smem_tc);
        if (e == cudaSuccess) {
            long long grid = n->num_sms;
            if (grid > B) grid = B;
            kern<<<(unsigned)grid, kEtThreads, n->smem_tc, (cudaStream_t)stream>>>(n->net_dev, x_dev, B, logits_dev, y2, y3, y4);
            e = cudaGetLastError();
            which = 1;
        }
    } else if (n->KS && !(force && force[0] == '1')) {
        which = 2;
        using kern_t = void (*)(const EegNet*, const int8_t*, long long, int8_t*, int8_t*, int8_t*, int8_t*);
        kern_t kern = nullptr;
#define MIBMI_EEG_PICK(d, ks) \
        if (n->D == d && n->KS == ks)                                                                                     \
            kern = n->F1 <= 8 ? (n->magic1 ? eegnet_mma_kernel<d, ks, true, 256> : eegnet_mma_kernel<d, ks, false, 256>)  \
                              : (n->magic1 ? eegnet_mma_kernel<d, ks, true, 512> : eegnet_mma_kernel<d, ks, false, 512>);
        MIBMI_EEG_PICK(1, 2) MIBMI_EEG_PICK(1, 3) MIBMI_EEG_PICK(1, 5) MIBMI_EEG_PICK(2, 2) MIBMI_EEG_PICK(2, 3) MIBMI_EEG_PICK(2, 5)
#undef MIBMI_EEG_PICK
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)n->smem_mma);
        if (e == cudaSuccess) {
            long long grid = (long long)n->num_sms * (n->F1 > 8 ? 1 : n->smem_mma > 72 * 1024 ? 2 : MIBMI_EEG_MINB);
            if (grid > B) grid = B;
            kern<<<(unsigned)grid, 32 * n->F1, n->smem_mma, (cudaStream_t)stream>>>(n->net_dev, x_dev, B, logits_dev, y2, y3, y4);
            e = cudaGetLastError();
        }
    } else {
        which = 3;
        auto kern = n->D == 1 ? (n->magic1 ? eegnet_forward_kernel<1, true> : eegnet_forward_kernel<1, false>)
                              : (n->magic1 ? eegnet_forward_kernel<2, true> : eegnet_forward_kernel<2, false>);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)n->smem);
        if (e == cudaSuccess) {
            long long grid = (long long)n->num_sms * (n->smem > 100 * 1024 ? 1 : 2);
            if (grid > B) grid = B;
            kern<<<(unsigned)grid, n->threads, n->smem, (cudaStream_t)stream>>>(n->net_dev, x_dev, B, logits_dev, y2, y3, y4);
            e = cudaGetLastError();
        }
    }
    if (e != cudaSuccess) {
        n->last_cuda.store((int)e);
        return MIBMI_ERR_CUDA;
    }
    n->launches.fetch_add(1);
    n->last_kernel.store(which);
    return MIBMI_OK;
}

const char* mibmi_eegnet_last_kernel(const mibmi_eegnet_t* n) {
    static const char* const names[] = {"", "eegnet_tc_kernel", "eegnet_mma_kernel", "eegnet_forward_kernel"};
    return n ? names[n->last_kernel.load() & 3] : "";
}

int mibmi_eegnet_forward_int8(mibmi_eegnet_t* n, const int8_t* x_dev, int64_t B, int8_t* logits_dev, void* stream) {
    return mibmi_eegnet_forward_int8_taps(n, x_dev, B, logits_dev, nullptr, nullptr, nullptr, stream);
}

int64_t mibmi_eegnet_launch_count(const mibmi_eegnet_t* n) { return n ? n->launches.load() : 0; }

}  // extern "C"

#ifdef MIBMI_PROF
extern "C" __attribute__((visibility("default"))) int mibmi_debug_prof_et(long long* out_host) {
    return (int)cudaMemcpyFromSymbol(out_host, mibmi::g_prof_et, sizeof(mibmi::g_prof_et));
}
#endif
