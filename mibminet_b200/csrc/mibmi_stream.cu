// Sliding-window streaming scoring for sm_100a (BASELINE config 5).
//
// Input: continuous int8 streams x[S subjects][C][L]; output: logits[S][W][4] of the W = (L - T) / 16 + 1
// windows of T samples starting every 16 samples -- bit-identical to running the network on every
// materialised window (net_model_compute, edge-eegnet_wolf/src/cl/net/model.c:42, on x[s][:, 16 w : 16 w + T]),
// including the per-window zero padding of layers 2 and 3 (gen_net_header.py:145-146, 171-172).
//
// A CTA (16 warps, two CTAs per SM) scores a chunk of G = 32 consecutive windows of one subject:
//   P0  the chunk's samples reach shared memory as 2-D TMA tiles [C][128] (cp.async.bulk.tensor, 128-byte
//       swizzle so that the transposing ldmatrix of layer 1 is bank-conflict free), half a chunk at a time
//   P1  layer 1 once per sample (not once per window): mma.sync m16n8k32.s8 as in mibmi_pipe.cu
//   P2  warp k = temporal channel k.  Layer 2 + ReLU threshold + sum-pool once per 8 stream samples
//       (Hankel x Toeplitz MMA); the pooled value is valid for every window in whose interior it lies.  It is
//       stored 8 times, shifted by 2 bytes each, so that the rows of 16 consecutive windows are all 16-byte
//       aligned for ldmatrix.  The 4 + 4 pooled samples at the two edges of every window, whose receptive
//       field is cut by the window's zero padding, are a masked-Toeplitz MMA with the WINDOW as the row
//   P3  warp k, 16 windows at a time: layer 3 with the window as the MMA row (masked interior + edge part)
//   P4  warp = window: layers 4 + 5 exactly as in the pipeline kernel
// Per window this executes ~1/6 of the instructions (and 1/6 of the IMMAs) of the per-trial kernel and reads
// 64 x 16 new bytes from HBM instead of 64 x 480.
#include <cuda.h>

#include "mibmi_common.cuh"
#include "mibmi_kernels.h"
#include "mibmi_ptx.cuh"
#include "mibmi_tmap.h"

namespace mibmi {

constexpr int kStreamWarps = 16;
constexpr int kStreamThreads = kStreamWarps * 32;

template <int C_, int T_>
struct SCfg {
    static constexpr int C = C_, T = T_, HOP = 16, G = 32;        // windows per chunk
    static constexpr int SPAN = T + HOP * (G - 1);                 // samples a chunk touches
    static constexpr int UNITS = (SPAN + 31) / 32;                 // layer-1 units of 32 samples
    static constexpr int HALF_UNITS = kStreamWarps;                // one unit per warp and half chunk
    static constexpr int HALF_TILES = HALF_UNITS / 4;              // TMA tiles of 128 samples per half
    static constexpr int CR = (C + 7) / 8 * 8, TILE_BYTES = CR * 128;
    static constexpr int KB16 = (C + 15) / 16, KP = KB16 / 2, KR = KB16 % 2;
    static constexpr int T8 = T / 8, T64 = T8 / 8;
    static constexpr int NCH = 36, PLANE = 16 * NCH, Y1S = 2 * PLANE + 16;   // PLANE = 64 mod 128, Y1S / 16 odd
    static constexpr int JMAX = 2 * (G - 1) + T8 - 5;              // last interior pooled sample of a chunk
    static constexpr int MT2 = JMAX / 64 + 1;                      // layer-2 m-tiles of 16 rows x 32 samples
    static constexpr int RCH = (T - 64) / 16 + 2;                  // y1 chunk of a window's last 64 samples
    static constexpr int CPS = 144, S2K = 8 * CPS;                 // shifted copies: stride = 16 mod 128
    static constexpr int NB3 = (T8 + 15) / 16;                     // layer-3 blocks of 16 outputs
    static constexpr int Y3S = 80;                                 // layer-3 row of a (window, channel)
    static constexpr int Y3W = 16 * Y3S + 32;                      // bytes per window; + 32: the 8 windows a store instruction touches
                                                                   // fall into 4 bank groups instead of one (was 8-way conflicted)
    static constexpr int NST = (T64 + 3) / 4, T64P = (T64 + 1) / 2 * 2, KST5 = T64P / 2;
    static constexpr int Y4_BYTES = ((16 * T64P + 15) / 16) * 16;
    // shared memory map
    static constexpr int BAR_A = 0, BAR_B = 8;
    static constexpr int OFF_X = 1024;
    static constexpr int OFF_Y1 = OFF_X + HALF_TILES * TILE_BYTES;
    static constexpr int OFF_S2 = OFF_Y1 + 16 * Y1S + 16;          // 16 B slack: copy 0 is written from byte -6
    static constexpr int OFF_E = OFF_S2 + 16 * S2K;
    static constexpr int OFF_Y3 = OFF_E + 16 * G * 16;
    static constexpr int OFF_Y4 = OFF_Y3 + 16 * Y3W;
    // the small constants every warp needs in every chunk, staged once per CTA (words).  With two CTAs of ~100 KB of shared memory per
    // SM the L1 is ~28 KB and the per-channel fragments streaming through it (147 KB per chunk) evict everything: these loads were L2
    // round trips (79 LDG per window, 13 % of the stall samples, ncu r1f)
    static constexpr int CW_W1 = 0, CW_OFF1 = 256, CW_M1 = 272, CW_SH1 = 288, CW_C2 = 304 /* thr, sig, soff, m, sh [16] */,
                         CW_W4 = 384, CW_W5 = 448, CW_C4 = 704 /* thr, sig, soff, m, sh [16] */, CW_MISC = 784 /* b5[4], m5, sh5, m3, sh3 */,
                         CW_WORDS = 792;
    static constexpr int OFF_CONS = ((OFF_Y4 + kStreamWarps * Y4_BYTES + 127) / 128) * 128;
    static constexpr int SMEM = ((OFF_CONS + 4 * CW_WORDS + 127) / 128) * 128;
    static_assert(KB16 <= 4, "w1nat block of the constant area");
    static_assert(T % 64 == 0 || T % 16 == 0, "window length");
    static_assert(UNITS <= 2 * HALF_UNITS, "a chunk is two half loads");
    static_assert(2 * (UNITS + 1) + 1 < 2 * NCH && 2 * (16 * MT2 - 1) + 5 < 2 * NCH, "y1 planes");
    static_assert(NB3 == 4 && T8 == 60, "layer-3 window form is written for T = 480");
    static_assert(KST5 <= 4, "classifier fragments live in registers");
    static_assert(TILE_BYTES % 1024 == 0, "swizzle atom");
    static_assert(SMEM <= 113 * 1024, "two CTAs per SM");
};

template <bool FAST>
__device__ __forceinline__ int sdiv1(int x, int m, int s) {
    if (FAST) return __mulhi(x, m) + (int)((unsigned)x >> 31);
    return (__mulhi(x, m) >> s) + (int)((unsigned)x >> 31);
}

// work item -> (subject, first window)
struct StreamItem { long long subj; int w0; };

template <class K, bool FAST1, bool TAPS, bool FAST3 = false>
__global__ void __launch_bounds__(kStreamThreads, 2)
stream_forward_kernel(const __grid_constant__ CUtensorMap tmx, const StreamNet* __restrict__ net, long long S, int W,
                      int chunks_per_subject, int8_t* __restrict__ logits, int8_t* __restrict__ tap2i,
                      int8_t* __restrict__ tap2e, int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const uint32_t sbase = smem_u32(smem);
    const long long nitems = S * chunks_per_subject;

    if (tid == 0) {
        mbar_init(sbase + K::BAR_A, 1);
        mbar_init(sbase + K::BAR_B, 1);
        mbar_fence_init();
    }
    {
        uint4* z = reinterpret_cast<uint4*>(smem + K::OFF_Y1);
        for (int i = tid; i < (K::OFF_CONS - K::OFF_Y1) / 16; i += kStreamThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    uint32_t* const cw = reinterpret_cast<uint32_t*>(smem + K::OFF_CONS);
    const int* const ci = reinterpret_cast<const int*>(cw);
    for (int i = tid; i < K::CW_WORDS; i += kStreamThreads) {
        uint32_t v = 0;
        if (i < K::CW_OFF1) v = (&net->w1nat[0][0][0])[i];
        else if (i < K::CW_M1) v = (uint32_t)net->off1[i - K::CW_OFF1];
        else if (i < K::CW_SH1) v = (uint32_t)net->m1[i - K::CW_M1];
        else if (i < K::CW_C2) v = (uint32_t)net->sh1[i - K::CW_SH1];
        else if (i < K::CW_W4) {
            const int f = (i - K::CW_C2) >> 4, kk = (i - K::CW_C2) & 15;
            v = (uint32_t)(f == 0 ? net->thr2[kk] : f == 1 ? net->sig2[kk] : f == 2 ? net->soff2[kk] : f == 3 ? net->m2[kk] : net->sh2[kk]);
        } else if (i < K::CW_W5) v = (&net->w4frag[0][0])[i - K::CW_W4];
        else if (i < K::CW_C4) v = (i - K::CW_W5) < K::KST5 * 64 ? (&net->w5frag[0][0][0])[i - K::CW_W5] : 0u;
        else if (i < K::CW_MISC) {
            const int f = (i - K::CW_C4) >> 4, kk = (i - K::CW_C4) & 15;
            v = (uint32_t)(f == 0 ? net->thr4[kk] : f == 1 ? net->sig4[kk] : f == 2 ? net->soff4[kk] : f == 3 ? net->m4[kk] : net->sh4[kk]);
        } else {
            const int j = i - K::CW_MISC;
            v = (uint32_t)(j < 4 ? net->b5[j] : j == 4 ? net->m5 : j == 5 ? net->sh5 : j == 6 ? net->m3 : net->sh3);
        }
        cw[i] = v;
    }
    __syncthreads();

    auto issue_half = [&](long long item, int half) {      // one thread
        const long long subj = item / chunks_per_subject;
        const int w0 = (int)(item % chunks_per_subject) * K::G;
        const uint32_t bar = sbase + (half ? K::BAR_B : K::BAR_A);
        mbar_expect_tx(bar, K::HALF_TILES * K::C * 128);
#pragma unroll
        for (int i = 0; i < K::HALF_TILES; i++)
            tma_load_2d(sbase + K::OFF_X + i * K::TILE_BYTES, &tmx, K::HOP * w0 + 128 * (K::HALF_TILES * half + i),
                        (int)(subj * K::C), bar);
    };
    if (tid == 0 && (long long)blockIdx.x < nitems) issue_half(blockIdx.x, 0);

    // ---- per-thread constants -------------------------------------------------------------------------
    // layer 1 (all warps)
    uint32_t rowoff[K::KP + 1], rowkey[K::KP + 1];
#pragma unroll
    for (int s = 0; s <= K::KP; s++) {
        int r = (s < K::KP) ? 16 * (2 * s + (lane >> 4)) + (lane & 15) : 16 * (K::KB16 - 1) + (lane & 15);
        if (r >= K::C) r &= 7;                         // zero weights; keeps the 8 rows of a phase distinct mod 8
        rowoff[s] = r * 128;
        rowkey[s] = (r & 7) << 4;
    }
    const uint32_t st_off = ((lane & 7) + 8 * ((lane >> 3) & 1)) * K::Y1S + (lane >> 4) * K::PLANE + 16;
    // layers 2 / 3: warp = channel
    const int k = warp;
    const int lrow = (lane & 7) + 8 * ((lane >> 3) & 1);
    const uint32_t y1k = sbase + K::OFF_Y1 + k * K::Y1S;
    const uint32_t l2_off = (lane >> 4) * K::PLANE + 16 * lrow;
    uint8_t* s2k = smem + K::OFF_S2 + k * K::S2K;
    uint8_t* ek = smem + K::OFF_E + k * (K::G * 16);
    // layer 4 / 5: warp = window
    const uint32_t l4_off = (lane & 15) * K::Y3S + 16 * (lane >> 4);
    uint8_t* y4 = smem + K::OFF_Y4 + warp * K::Y4_BYTES;
    const uint32_t* y4w = reinterpret_cast<const uint32_t*>(y4);

    int it = 0;
    for (long long item = blockIdx.x; item < nitems; item += gridDim.x, it++) {
        const long long subj = item / chunks_per_subject;
        const int w0 = (int)(item % chunks_per_subject) * K::G;

        // ================================================================================================
        // P1: layer 1 over the chunk's samples, half a chunk (16 units of 32 samples) at a time
        // ================================================================================================
#pragma unroll 1
        for (int half = 0; half < 2; half++) {
            mbar_wait(sbase + (half ? K::BAR_B : K::BAR_A), it & 1);
            const int u = K::HALF_UNITS * half + warp;
            if (u < K::UNITS) {
                uint32_t wfr[K::KB16][2];
#pragma unroll
                for (int kb = 0; kb < K::KB16; kb++) {
                    wfr[kb][0] = cw[K::CW_W1 + (2 * kb) * 32 + lane];
                    wfr[kb][1] = cw[K::CW_W1 + (2 * kb + 1) * 32 + lane];
                }
                const uint32_t xt = sbase + K::OFF_X + (warp >> 2) * K::TILE_BYTES;
                const uint32_t j0 = 2 * (warp & 3);
                int acc[2][2][4];
#pragma unroll
                for (int nt = 0; nt < 2; nt++) {
                    const int o0 = ci[K::CW_OFF1 + 8 * nt + 2 * q], o1 = ci[K::CW_OFF1 + 8 * nt + 2 * q + 1];
#pragma unroll
                    for (int mt = 0; mt < 2; mt++) {
                        acc[mt][nt][0] = acc[mt][nt][2] = o0;
                        acc[mt][nt][1] = acc[mt][nt][3] = o1;
                    }
                }
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    const uint32_t jc = (j0 + mt) << 4;
#pragma unroll
                    for (int s = 0; s < K::KP; s++) {
                        uint32_t d[4];
                        ldsm_t_x2(d, xt + rowoff[s] + (jc ^ rowkey[s]));
                        mma_k32(acc[mt][0], d, wfr[2 * s][0], wfr[2 * s + 1][0]);
                        mma_k32(acc[mt][1], d, wfr[2 * s][1], wfr[2 * s + 1][1]);
                    }
                    if (K::KR) {
                        uint32_t d[2];
                        ldsm_t_x1(d, xt + rowoff[K::KP] + (jc ^ rowkey[K::KP]));
                        mma_k16(acc[mt][0], d[0], d[1], wfr[K::KB16 - 1][0]);
                        mma_k16(acc[mt][1], d[0], d[1], wfr[K::KB16 - 1][1]);
                    }
                }
                uint32_t w[2][2];
#pragma unroll
                for (int nt = 0; nt < 2; nt++) {
                    const int mg0 = ci[K::CW_M1 + 8 * nt + 2 * q], mg1 = ci[K::CW_M1 + 8 * nt + 2 * q + 1];
                    const int s0 = ci[K::CW_SH1 + 8 * nt + 2 * q], s1 = ci[K::CW_SH1 + 8 * nt + 2 * q + 1];
#pragma unroll
                    for (int mt = 0; mt < 2; mt++) {
                        const int v0 = sdiv1<FAST1>(acc[mt][nt][0], mg0, s0), v1 = sdiv1<FAST1>(acc[mt][nt][1], mg1, s1);
                        const int v2 = sdiv1<FAST1>(acc[mt][nt][2], mg0, s0), v3 = sdiv1<FAST1>(acc[mt][nt][3], mg1, s1);
                        w[mt][nt] = pack4_sat(v0, v1, v2, v3);
                    }
                }
                stsm_t8_x4(sbase + K::OFF_Y1 + st_off + 16 * u, w[0][0], w[0][1], w[1][0], w[1][1]);
            }
            __syncthreads();      // the x buffer is free, this half of y1 is visible
            if (tid == 0) {
                fence_proxy_async();
                if (half == 0) issue_half(item, 1);
                else if (item + gridDim.x < nitems) issue_half(item + gridDim.x, 0);
            }
        }

        // ================================================================================================
        // P2: warp k = channel k.  (a) stream-level layer 2 + pooling, stored as 8 shifted copies
        // ================================================================================================
        {
            const int thr = ci[K::CW_C2 + k], sig = ci[K::CW_C2 + 16 + k], soff = ci[K::CW_C2 + 32 + k];
            const int m2 = ci[K::CW_C2 + 48 + k], sh2 = ci[K::CW_C2 + 64 + k];
            {
                uint32_t bt[4][3][2];
#pragma unroll
                for (int v = 0; v < 4; v++)
#pragma unroll
                    for (int s = 0; s < 3; s++)
#pragma unroll
                        for (int h = 0; h < 2; h++) bt[v][s][h] = __ldg(&net->w2frag[k][v * 6 + s * 2 + h][lane]);
#pragma unroll 1
                for (int mt = 0; mt < K::MT2; mt++) {
                    int acc[4][4];
#pragma unroll
                    for (int v = 0; v < 4; v++) acc[v][0] = acc[v][1] = acc[v][2] = acc[v][3] = 0;
#pragma unroll
                    for (int s = 0; s < 3; s++) {
                        uint32_t a[4];
                        ldsm_x4(a, y1k + l2_off + 16 * (16 * mt + s));
#pragma unroll
                        for (int v = 0; v < 4; v++) mma_k32(acc[v], a, bt[v][s][0], bt[v][s][1]);
                    }
#pragma unroll
                    for (int hf = 0; hf < 2; hf++) {
                        int sum = 0;
#pragma unroll
                        for (int v = 0; v < 4; v++) sum += max(acc[v][2 * hf], thr) + max(acc[v][2 * hf + 1], thr);
                        const int j = 4 * (16 * mt + g + 8 * hf) + q;       // pooled stream sample of the chunk
                        const int o = clip8(tdiv_magic(sum * sig + soff, m2, sh2));
                        // copy r holds sample j at byte j + 8 - 2 r: the row of window 8 m + r (its samples
                        // p = 16 b - 8 ..) starts at byte 16 (m + b).  Bytes -6 .. -1 / 128 .. 135 fall into padding.
                        uint8_t* d = s2k + j + 8;
#pragma unroll
                        for (int r = 0; r < 8; r++) d[r * (K::CPS - 2)] = (uint8_t)o;
                        if (TAPS && tap2i) {
                            const long long js = 2LL * w0 + j;              // pooled sample of the stream
                            if (j >= 4 && j <= K::JMAX && js <= (long long)2 * (W - 1) + K::T8 - 5) tap2i[((size_t)subj * 16 + k) * (2 * (size_t)(W - 1) + K::T8) + js] = (int8_t)o;
                        }
                    }
                }
            }
            // (b) the 4 + 4 pooled samples at the edges of every window: masked Toeplitz, window = MMA row
#pragma unroll 1
            for (int side = 0; side < 2; side++) {
                uint32_t be[4][2][2];
                const uint32_t(*src)[32] = side ? net->w2R[k] : net->w2L[k];
#pragma unroll
                for (int v = 0; v < 4; v++)
#pragma unroll
                    for (int s = 0; s < 2; s++)
#pragma unroll
                        for (int h = 0; h < 2; h++) be[v][s][h] = __ldg(&src[v * 4 + s * 2 + h][lane]);
#pragma unroll 1
                for (int wt = 0; wt < K::G / 16; wt++) {
                    int acc[4][4];
#pragma unroll
                    for (int v = 0; v < 4; v++) acc[v][0] = acc[v][1] = acc[v][2] = acc[v][3] = 0;
#pragma unroll
                    for (int s = 0; s < 2; s++) {
                        // row = window 16 wt + lrow; its 16-byte chunk c lives in plane c & 1 at index c >> 1
                        const int c = 16 * wt + lrow + (side ? K::RCH : 2) + 2 * s + (lane >> 4);
                        uint32_t a[4];
                        ldsm_x4(a, y1k + (c & 1) * K::PLANE + 16 * (c >> 1));
#pragma unroll
                        for (int v = 0; v < 4; v++) mma_k32(acc[v], a, be[v][s][0], be[v][s][1]);
                    }
#pragma unroll
                    for (int hf = 0; hf < 2; hf++) {
                        int sum = 0;
#pragma unroll
                        for (int v = 0; v < 4; v++) sum += max(acc[v][2 * hf], thr) + max(acc[v][2 * hf + 1], thr);
                        const int o = clip8(tdiv_magic(sum * sig + soff, m2, sh2));
                        const int wl = 16 * wt + g + 8 * hf;
                        ek[wl * 16 + 4 * side + q] = (uint8_t)o;
                        if (TAPS && tap2e && w0 + wl < W)
                            tap2e[(((size_t)subj * W + w0 + wl) * 16 + k) * 8 + 4 * side + q] = (int8_t)o;
                    }
                }
            }
        }
        __syncthreads();

        // ================================================================================================
        // P3 / P4, 16 windows at a time
        // ================================================================================================
#pragma unroll 1
        for (int wt = 0; wt < K::G / 16; wt++) {
            // ---- P3: warp k, layer 3 of windows 16 wt .. + 15; D[window][16 outputs] per block b
            {
                const int m3 = ci[K::CW_MISC + 6], sh3 = ci[K::CW_MISC + 7];
                const int gw = 16 * wt + lrow;                                  // this lane's ldmatrix row
                const uint32_t arow = smem_u32(s2k) + (gw & 7) * K::CPS + 16 * (gw >> 3) + 16 * (lane >> 4);
                uint32_t ae[2];
                ldsm_x2(ae, smem_u32(ek) + (16 * wt + (lane & 15)) * 16);
                uint8_t* y3k = smem + K::OFF_Y3 + k * K::Y3S + 16 * (q >> 1) + 8 * (q & 1);
#pragma unroll
                for (int b = 0; b < K::NB3; b++) {
                    const int cls = b == 0 ? 0 : (b == K::NB3 - 1 ? 2 : 1);
                    uint32_t a[4];
                    ldsm_x4(a, arow + 16 * b);
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        int c3[4] = {0, 0, 0, 0};
                        mma_k32(c3, a, __ldg(&net->w3s[k][cls * 4 + e * 2][lane]), __ldg(&net->w3s[k][cls * 4 + e * 2 + 1][lane]));
                        if (b == 0 || b == K::NB3 - 1)
                            mma_k16(c3, ae[0], ae[1], __ldg(&net->w3e[k][(b == 0 ? 0 : 2) + e][lane]));
                        const int t = 16 * b + 8 * e + 2 * q;
                        if (t < K::T8) {
                            // the sample order layer 4 wants (mibmi_pipe.cu): pooling group = MMA column pair
                            const int pos = 32 * (b >> 1) + 4 * (b & 1) + 2 * e;
#pragma unroll
                            for (int hf = 0; hf < 2; hf++) {
                                const int wl = g + 8 * hf;
                                const uint32_t w = FAST3 ? pack2_sat(sdiv1<true>(c3[2 * hf], m3, 0), sdiv1<true>(c3[2 * hf + 1], m3, 0))
                                                         : pack2_sat(tdiv_magic(c3[2 * hf], m3, sh3), tdiv_magic(c3[2 * hf + 1], m3, sh3));
                                *reinterpret_cast<uint16_t*>(y3k + wl * K::Y3W + pos) = (uint16_t)w;
                                if (TAPS && tap3 && w0 + 16 * wt + wl < W) {
                                    int8_t* o = tap3 + (((size_t)subj * W + w0 + 16 * wt + wl) * 16 + k) * K::T8 + t;
                                    o[0] = (int8_t)(w & 0xff);
                                    o[1] = (int8_t)((w >> 8) & 0xff);
                                }
                            }
                        }
                    }
                }
            }
            __syncthreads();
            // ---- P4: warp = window 16 wt + warp, layers 4 + 5
            {
                const int wl = warp, wg = w0 + 16 * wt + wl;
                const uint32_t w4a0 = cw[K::CW_W4 + lane], w4a1 = cw[K::CW_W4 + 32 + lane];
                const uint32_t y3a = sbase + K::OFF_Y3 + wl * K::Y3W + l4_off;
#pragma unroll
                for (int s4 = 0; s4 < K::NST; s4++) {
                    uint32_t d[4];
                    ldsm_t_x2(d, y3a + 32 * s4);
                    int c[4][4];
#pragma unroll
                    for (int v = 0; v < 4; v++) {
                        c[v][0] = c[v][1] = c[v][2] = c[v][3] = 0;
                        mma_k16(c[v], w4a0, w4a1, d[v]);
                    }
                    const int thr_lo = ci[K::CW_C4 + g], thr_hi = ci[K::CW_C4 + g + 8];
                    int s_lo = 0, s_hi = 0;
#pragma unroll
                    for (int v = 0; v < 4; v++) {
                        s_lo += max(c[v][0], thr_lo) + max(c[v][1], thr_lo);
                        s_hi += max(c[v][2], thr_hi) + max(c[v][3], thr_hi);
                    }
                    const int grp = 4 * s4 + q;
                    const int o_lo = clip8(tdiv_magic(s_lo * ci[K::CW_C4 + 16 + g] + ci[K::CW_C4 + 32 + g], ci[K::CW_C4 + 48 + g], ci[K::CW_C4 + 64 + g]));
                    const int o_hi = clip8(tdiv_magic(s_hi * ci[K::CW_C4 + 24 + g] + ci[K::CW_C4 + 40 + g], ci[K::CW_C4 + 56 + g], ci[K::CW_C4 + 72 + g]));
                    if (grp < K::T64) {
                        y4[g * K::T64P + grp] = (uint8_t)o_lo;
                        y4[(g + 8) * K::T64P + grp] = (uint8_t)o_hi;
                        if (TAPS && tap4 && wg < W) {
                            tap4[(((size_t)subj * W + wg) * 16 + g) * K::T64 + grp] = (int8_t)o_lo;
                            tap4[(((size_t)subj * W + wg) * 16 + g + 8) * K::T64 + grp] = (int8_t)o_hi;
                        }
                    }
                }
                __syncwarp();
                int z[4] = {0, 0, 0, 0};
#pragma unroll
                for (int ks = 0; ks < K::KST5; ks++) {
                    const uint32_t a[4] = {cw[K::CW_W5 + (2 * ks) * 32 + lane], 0u, cw[K::CW_W5 + (2 * ks + 1) * 32 + lane], 0u};
                    const uint32_t b0 = g == 0 ? y4w[8 * ks + q] : 0u, b1 = g == 0 ? y4w[8 * ks + 4 + q] : 0u;
                    mma_k32(z, a, b0, b1);
                }
                if (q == 0 && g < 4 && wg < W)
                    logits[((size_t)subj * W + wg) * 4 + g] =
                        (int8_t)clip8(tdiv_magic(z[0] + ci[K::CW_MISC + (g & 3)], ci[K::CW_MISC + 4], ci[K::CW_MISC + 5]));
                __syncwarp();
            }
            __syncthreads();      // y3 (and, after the last tile, y1 / the copies / the edges) may be rewritten
        }
    }
}

// ---------------------------------------------------------------------------------------------
template <int C, int T>
static cudaError_t launch_stream_instance(const StreamNet* net_dev, int fast, const int8_t* x, long long S, long long L,
                                          int8_t* logits, int8_t* t2i, int8_t* t2e, int8_t* t3, int8_t* t4, int num_sms,
                                          cudaStream_t stream) {
    using K = SCfg<C, T>;
    tmap_encode_t enc = tmap_encoder();
    if (!enc) return cudaErrorNotSupported;
    const long long W = (L - T) / K::HOP + 1;
    const int chunks = (int)((W + K::G - 1) / K::G);
    CUtensorMap tm;
    cuuint64_t dims[2] = {(cuuint64_t)L, (cuuint64_t)(S * C)};
    cuuint64_t strides[1] = {(cuuint64_t)L};
    cuuint32_t box[2] = {128, (cuuint32_t)C};
    cuuint32_t estr[2] = {1, 1};
    if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<int8_t*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorInvalidValue;
    const bool taps = t2i || t2e || t3 || t4;
    auto kern = taps ? stream_forward_kernel<K, false, true>
                     : ((fast & 1) ? ((fast & 2) ? stream_forward_kernel<K, true, false, true> : stream_forward_kernel<K, true, false>)
                                   : stream_forward_kernel<K, false, false>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
    if (e != cudaSuccess) return e;
    long long grid = 2LL * num_sms;
    if (grid > S * chunks) grid = S * chunks;
    kern<<<(unsigned)grid, kStreamThreads, K::SMEM, stream>>>(tm, net_dev, S, (int)W, chunks, logits, t2i, t2e, t3, t4);
    return cudaGetLastError();
}

#define MIBMI_STREAM_SHAPES(X) X(64, 480) X(38, 480) X(24, 480) X(19, 480) X(16, 480) X(8, 480)

bool stream_has_instance(int C, int T) {
#define X(c, t) if (C == c && T == t) return true;
    MIBMI_STREAM_SHAPES(X)
#undef X
    return false;
}

cudaError_t launch_stream(const StreamNet* net_dev, int fast, int C, int T, const int8_t* x, long long S, long long L,
                          int8_t* logits, int8_t* t2i, int8_t* t2e, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream) {
#define X(c, t) if (C == c && T == t) return launch_stream_instance<c, t>(net_dev, fast, x, S, L, logits, t2i, t2e, t3, t4, num_sms, stream);
    MIBMI_STREAM_SHAPES(X)
#undef X
    return cudaErrorInvalidValue;
}

// Fallback for any hop / shape: materialise windows [nw][C][T] from the streams (then run the per-trial path)
__global__ void gather_windows_kernel(const int8_t* __restrict__ x, long long L, int C, int T, int hop, long long W,
                                      long long first, long long nw, int8_t* __restrict__ out) {
    // one CTA per (window, channel) row; 16-byte vectors when everything is aligned
    const long long row = blockIdx.x;            // window-local index * C + channel
    const long long wi = first + row / C;
    const int c = (int)(row % C);
    const long long subj = wi / W, w = wi % W;
    const int8_t* src = x + ((size_t)subj * C + c) * L + w * hop;
    int8_t* dst = out + (size_t)row * T;
    if ((((uintptr_t)src | (uintptr_t)dst) & 15) == 0 && T % 16 == 0) {
        for (int i = threadIdx.x; i < T / 16; i += blockDim.x)
            reinterpret_cast<uint4*>(dst)[i] = __ldg(reinterpret_cast<const uint4*>(src) + i);
    } else {
        for (int i = threadIdx.x; i < T; i += blockDim.x) dst[i] = src[i];
    }
}

cudaError_t launch_gather_windows(const int8_t* x, long long L, int C, int T, int hop, long long W, long long first,
                                  long long nw, int8_t* out, cudaStream_t stream) {
    if (nw <= 0) return cudaSuccess;
    gather_windows_kernel<<<(unsigned)(nw * C), 64, 0, stream>>>(x, L, C, T, hop, W, first, nw, out);
    return cudaGetLastError();
}

}  // namespace mibmi
