// Warp-specialised int8 edgeEEGNet / MI-BMInet forward for sm_100a (the hot kernel of libmibmi.so).
//
// One persistent CTA per SM (25 warps) streams trials b = blockIdx.x + n * gridDim.x through all five
// layers; inter-layer activations never leave shared memory, the network never enters it (every warp
// keeps its MMA fragments of the weights in registers for the whole launch).
//
//   warp 24 "C"   TMA producer: two cp.async.bulk copies per trial into an S-stage ring (mbarrier
//                 complete_tx), then layers 4 + 5 of the trial whose layer 3 has finished
//   warps 0-7 "A"   layer 1, spatial conv: D[16 filters x 8 samples] += W1[16 x 32 ch] * X[32 ch x 8]
//                 with mma.sync m16n8k32.s8; the [C][T] -> channel-contiguous B operand comes out of
//                 ldmatrix.m16n16.trans.b8 (the two channel halves of a trial sit 16 B apart modulo
//                 128, so every 8-row phase is bank-conflict free); requant = one IMAD.HI (+ shift) +
//                 sign fix-up, saturating pack, one stmatrix per 32 samples
//                 -- edge-eegnet_wolf/src/cl/net/layer1.c:87-96
//   warps 8-23 "B"  one temporal channel each.  Layer 2, depthwise 64-tap conv as a Hankel x Toeplitz
//                 product: A[16 windows x 32] straight from the layer-1 row by ldmatrix (windows 32
//                 samples apart; even / odd 16-byte chunks live in two planes so that a phase reads 128
//                 contiguous bytes), B = 4 Toeplitz variants of the taps held in 24 registers; every
//                 thread ends up with complete pooling groups of 8, so ReLU-threshold + sum-pool +
//                 requant need no shuffles -- layer2.c:87-118 (REORDER_BN).
//                 Layer 3, depthwise 16-tap conv, the same trick with one k-step -- layer3.c:70
//   warp 24       layer 4, pointwise conv: mma.sync m16n8k16 over a sample order that layer 3 wrote so
//                 that a thread again owns whole pooling groups -- layer4.c:99-140; layer 5, classifier,
//                 4 MMAs against the [16][T64] activations -- layer5.c:81-89
//
// Hand-off between the roles is mbarrier-only (no __syncthreads after the prologue):
//   full_x / empty_x [S]   producer <-> A      full_y1 / empty_y1 [2]   A <-> B
//   full_y3 / empty_y3 [2] B <-> C
// Integer results are bit-identical to the reference (tests/test_parity_gpu.py).
#include <cuda.h>
#include <string.h>

#include "mibmi_common.cuh"
#include "mibmi_kernels.h"
#include "mibmi_ptx.cuh"
#include "mibmi_tmap.h"

namespace mibmi {

#ifdef MIBMI_PROF
// experiment build only (nvcc -DMIBMI_PROF): per-warp cycles spent in each mbarrier wait, read by tools/prof_pipe.py
__device__ long long g_prof[148 * 32 * 4];
#define PWAIT(slot, ...) do { if (!mbar_test(__VA_ARGS__)) { const long long t0_ = clock64(); mbar_wait(__VA_ARGS__); pw_[slot] += clock64() - t0_; } } while (0)
#define PROF_BEGIN long long pw_[3] = {0, 0, 0}; const long long pt0_ = clock64();
#define PROF_END if ((threadIdx.x & 31) == 0) { long long* o_ = g_prof + (blockIdx.x * 32 + (threadIdx.x >> 5)) * 4; \
    o_[0] = clock64() - pt0_; o_[1] = pw_[0]; o_[2] = pw_[1]; o_[3] = pw_[2]; }
#else
#define PWAIT(slot, ...) mbar_wait(__VA_ARGS__)
#define PROF_BEGIN
#define PROF_END
#endif

constexpr int kAW = 8;                        // layer-1 warps
constexpr int kBW = 16;                       // layer-2/3 warps (one channel each)
constexpr int kCW = 2;                         // layer-4/5 warps (alternate trials)
constexpr int kPipeThreads = (kAW + kBW + kCW + 2) * 32;   // + TMA producer warp + tcgen05.mma issuer warp
#ifndef MIBMI_UMMA
#define MIBMI_UMMA 1
#endif
#ifndef MIBMI_ROTATE
#define MIBMI_ROTATE 1
#endif
// experiment switches (tools/build_variant.py): role timings in isolation; results are wrong when set
#ifdef MIBMI_EXP_SKIP_A
#define EXP_SKIP_A && false
#else
#define EXP_SKIP_A
#endif
#ifdef MIBMI_EXP_SKIP_L2
#define EXP_SKIP_L2 && false
#else
#define EXP_SKIP_L2
#endif
#ifdef MIBMI_EXP_SKIP_L3
#define EXP_DO_L3 false
#else
#define EXP_DO_L3 true
#endif
#ifndef MIBMI_STAGES
#define MIBMI_STAGES 4
#endif
#ifndef MIBMI_NY1
#define MIBMI_NY1 3
#endif
#ifndef MIBMI_NY3
#define MIBMI_NY3 4
#endif
constexpr int kNY1 = MIBMI_NY1;                // layer-1 output ring (A -> B)
constexpr int kNY3 = MIBMI_NY3;                // layer-3 output ring (B -> C)

template <int C_, int T_, int S_>
struct PCfg {
    static constexpr int C = C_, T = T_, S = S_;
    static constexpr int KB16 = (C + 15) / 16, KP = KB16 / 2, KR = KB16 % 2;
    static constexpr int CA = (C + 1) / 2, CB = C - CA;          // channel halves (two bulk copies)
    // REAL: rows of T bytes are not 16-byte aligned in global memory (T % 16 != 0, e.g. 22 x 1125): every row
    // is bulk-copied from the 16-byte boundary below it and the layer-1 warps shift it into place (see A)
    static constexpr bool REAL = (T % 16) != 0;
    // UMMA: layer 1 on the 5th-generation tensor core.  The trial arrives as 2-D TMA tiles [KPAD ch][128 samples]
    // with the 128-byte swizzle = the canonical MN-major operand of tcgen05.mma (kind::i8, M = 128 samples, N = 16
    // filters, K = 32 channels per instruction); the int32 accumulators of a trial live in TMEM (16 columns per
    // tile, two trials in flight) and the layer-1 warps read them back with tcgen05.ld.16x256b, whose register
    // layout is the m16n8 fragment the requantisation epilogue was written for.
    static constexpr bool UMMA = MIBMI_UMMA && !REAL;
    static constexpr int TILES = (T + 127) / 128, KPAD = (C + 31) / 32 * 32, KSTEPS = KPAD / 32;
    static constexpr int TILE_BYTES = KPAD * 128;
    static constexpr int TMEM_COLS_TRIAL = 16 * TILES;            // accumulator columns of one trial
    static constexpr int TMEM_COLS = 2 * TMEM_COLS_TRIAL <= 32 ? 32 : 2 * TMEM_COLS_TRIAL <= 64 ? 64 :
                                     2 * TMEM_COLS_TRIAL <= 128 ? 128 : 2 * TMEM_COLS_TRIAL <= 256 ? 256 : 512;
    static constexpr int T8 = T / 8, T64 = T8 / 8;
    static constexpr int UNITS = (T + 31) / 32;                  // layer-1 units of 32 samples
    static constexpr int XP = REAL ? 32 * UNITS + 48 : T;        // row pitch inside a stage
    static constexpr int ROFF = REAL ? 0 : CA * T + 16;          // second half: +16 B modulo 128
    static constexpr int SCR_ROW = 48, SCR_WARP = 32 * SCR_ROW;  // REAL: per-warp [32 ch][32 samples (+16 pad)]
    static constexpr int MROWS = (T8 * 8 + 31) / 32, MT = (MROWS + 15) / 16;
    static constexpr int NCH = 16 * MT + 2;                      // 16-byte chunks per plane
    static constexpr int PLANE = 16 * NCH, Y1S = 2 * PLANE + 16; // Y1S / 16 odd: conflict-free stmatrix
    static constexpr int NW3 = (T8 + 15) / 16, Y2S = 16 * (NW3 + 1);
    static constexpr int Y3CH = 2 * ((NW3 + 1) / 2), Y3S = 16 * (Y3CH + 1);
    static constexpr int NST = (T64 + 3) / 4;                    // layer-4 super-tiles of 4 pooling groups
    static constexpr int T64P = (T64 + 1) / 2 * 2, KST5 = T64P / 2;
    static constexpr int TRIAL_BYTES = C * T;
    static constexpr int XS_BYTES = UMMA ? TILES * TILE_BYTES : (((REAL ? C * XP : C * T + 16) + 127) / 128) * 128;
    static constexpr int Y1_BYTES = 16 * Y1S, Y3_BYTES = 16 * Y3S;
    // barriers (8 B each)
    static constexpr int BAR_FULL_X = 0, BAR_EMPTY_X = 8 * S, BAR_FULL_Y1 = 16 * S, BAR_EMPTY_Y1 = BAR_FULL_Y1 + 8 * kNY1,
                         BAR_FULL_Y3 = BAR_EMPTY_Y1 + 8 * kNY1, BAR_EMPTY_Y3 = BAR_FULL_Y3 + 8 * kNY3,
                         BAR_TFULL = BAR_EMPTY_Y3 + 8 * kNY3, BAR_TEMPTY = BAR_TFULL + 16, BAR_END = BAR_TEMPTY + 16;
    // shared memory map
    static constexpr int OFF_TMEM = 248;                         // TMEM base address written by tcgen05.alloc
    static constexpr int OFF_W1 = 1024;                          // UMMA: W1 as 8 x 16-byte core matrices (K-major)
    static constexpr int W1_SBO = KSTEPS * 256;                  // bytes between the two groups of 8 filters
    static constexpr int OFF_XS = UMMA ? 3072 : 256;             // UMMA tiles: 1024-byte aligned swizzle atoms
    static constexpr int OFF_Y1 = OFF_XS + S * XS_BYTES;
    static constexpr int OFF_Y2 = OFF_Y1 + kNY1 * Y1_BYTES;
    static constexpr int OFF_Y3 = OFF_Y2 + 16 * Y2S;
    static constexpr int OFF_Y4 = OFF_Y3 + kNY3 * Y3_BYTES;
    static constexpr int Y4_BYTES = ((16 * T64P + 15) / 16) * 16;
    static constexpr int OFF_SCR = ((OFF_Y4 + kCW * Y4_BYTES + 15) / 16) * 16;
    static constexpr int SMEM_RAW = ((OFF_SCR + (REAL ? kAW * SCR_WARP : 0) + 127) / 128) * 128;
    static constexpr int SMEM = SMEM_RAW <= 227 * 1024 ? SMEM_RAW : 0;   // 0: this depth is never instantiated
    static_assert(REAL || T % 32 == 0, "aligned loader: whole 32-sample units");
    static_assert(!REAL || C <= 32, "realigning loader: one lane per channel");
    static_assert(T8 % 2 == 0, "layer 3 stores sample pairs");
    static_assert(KB16 >= 1 && KB16 <= kPipeMaxKB16, "C <= 64");
    static_assert(UNITS + 1 <= NCH - 1, "layer-1 chunks fit the planes with one zero chunk on each side");
    static_assert(BAR_END <= OFF_TMEM, "barrier area");
    static_assert(!UMMA || (2 * TMEM_COLS_TRIAL <= 512 && 16 * KPAD <= 2048), "TMEM / W1 staging");
    static_assert(KST5 <= kPipeMaxKst5, "T too large for PipeNet::w5frag");
};

// smem byte offset (within a stage) of row r of 16-channel block kb of the layer-1 B operand.
// Rows alternate between the two channel halves in groups of 4 so that any 8 consecutive rows hit 8
// different 16-byte bank groups (T / 16 = 2 mod 4 for T = 480); rows past C point at row 0 (zero weights).
template <class K>
__device__ __forceinline__ uint32_t l1_row_offset(int kb, int r) {
    const int region = (r >> 2) & 1;
    const int idx = 8 * kb + 4 * (r >> 3) + (r & 3);
    const bool valid = idx < (region ? K::CB : K::CA);
    return valid ? (uint32_t)(region ? K::ROFF + idx * K::T : idx * K::T) : 0u;
}

template <bool FAST>
__device__ __forceinline__ int tdiv1(int x, int m, int s) {
    if (FAST) return __mulhi(x, m) + (int)((unsigned)x >> 31);
    return (__mulhi(x, m) >> s) + (int)((unsigned)x >> 31);
}

template <class K, bool FAST1, bool TAPS>
__global__ void __launch_bounds__(kPipeThreads, 1)
pipe_forward_kernel(const __grid_constant__ CUtensorMap tmx, const PipeNet* __restrict__ net, const int8_t* __restrict__ x, long long B,
                    int8_t* __restrict__ logits, const PeerOut peers, int8_t* __restrict__ tap1, int8_t* __restrict__ tap2,
                    int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const uint32_t sbase = smem_u32(smem);
    const long long first = blockIdx.x, stride = gridDim.x;
    const int nloc = first < B ? (int)((B - first + stride - 1) / stride) : 0;

    // ---- prologue: barriers, zero the activation buffers (their padding stays zero for the whole launch)
    if (tid == 0) {
        for (int s = 0; s < K::S; s++) {
            mbar_init(sbase + K::BAR_FULL_X + 8 * s, 1);
            mbar_init(sbase + K::BAR_EMPTY_X + 8 * s, K::UMMA ? 1 : kAW);   // UMMA: one tcgen05.commit frees a stage
        }
        for (int i = 0; i < kNY1; i++) {
            mbar_init(sbase + K::BAR_FULL_Y1 + 8 * i, kAW);
            mbar_init(sbase + K::BAR_EMPTY_Y1 + 8 * i, kBW);
        }
        for (int i = 0; i < kNY3; i++) {
            mbar_init(sbase + K::BAR_FULL_Y3 + 8 * i, kBW);
            mbar_init(sbase + K::BAR_EMPTY_Y3 + 8 * i, 1);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(sbase + K::BAR_TFULL + 8 * i, 1);
            mbar_init(sbase + K::BAR_TEMPTY + 8 * i, kAW);
        }
        mbar_fence_init();
    }
    if (K::UMMA) {
        // W1 as the K-major B operand; TMEM for two trials' accumulators (allocated by the issuer warp)
        for (int i = tid; i < 16 * K::KPAD / 4; i += kPipeThreads)
            reinterpret_cast<uint32_t*>(smem + K::OFF_W1)[i] = reinterpret_cast<const uint32_t*>(net->w1umma)[i];
        if (warp == kAW + kBW + kCW + 1) tmem_alloc(sbase + K::OFF_TMEM, K::TMEM_COLS);
        fence_proxy_async();
        tc_fence_before();
    }
    {
        uint4* z = reinterpret_cast<uint4*>(smem + K::OFF_Y1);
        for (int i = tid; i < (K::SMEM - K::OFF_Y1) / 16; i += kPipeThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = K::UMMA ? *reinterpret_cast<const uint32_t*>(smem + K::OFF_TMEM) : 0u;
    PROF_BEGIN

    if (warp < kAW) {
        // =====================================================================================
        // A: layer 1
        // =====================================================================================
      if constexpr (K::UMMA) {
        // the contraction ran on the tensor core (issuer warp below); this warp owns TMEM lanes 32 (warp & 3) ..
        // + 31 = samples 32 (warp & 3) .. of every 128-sample tile, and tiles jb, jb + 2 (jb alternates per trial)
        int off1[2][2], mg1[2][2], sh1[2][2];
#pragma unroll
        for (int nt = 0; nt < 2; nt++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                off1[nt][e] = net->off1[8 * nt + 2 * q + e];
                mg1[nt][e] = net->m1[8 * nt + 2 * q + e];
                sh1[nt][e] = net->sh1[8 * nt + 2 * q + e];
            }
        const uint32_t st_off = ((lane & 7) + 8 * ((lane >> 3) & 1)) * K::Y1S + (lane >> 4) * K::PLANE + 16;
        const int qi = warp & 3;
        for (int n = 0; n < nloc; n++) {
            const int buf = n % kNY1, ab = n & 1;
            PWAIT(0, sbase + K::BAR_TFULL + 8 * ab, (n >> 1) & 1);
            tc_fence_after();
            PWAIT(1, sbase + K::BAR_EMPTY_Y1 + 8 * buf, ((n / kNY1) & 1) ^ 1);
            const uint32_t yb = sbase + K::OFF_Y1 + buf * K::Y1_BYTES + st_off;
#pragma unroll 1
            for (int j = ((warp >> 2) + (MIBMI_ROTATE ? n : 0)) & 1; j < K::TILES EXP_SKIP_A; j += 2) {
                const int u = 4 * j + qi;                        // unit of 32 samples
                if (32 * u >= K::T) break;
                uint32_t r[2][8];
                const uint32_t ta = tmem_base + ((uint32_t)(32 * qi) << 16) + ab * K::TMEM_COLS_TRIAL + 16 * j;
                tmem_ld_16x256b_x2(r[0], ta);
                tmem_ld_16x256b_x2(r[1], ta + (16u << 16));
                tmem_ld_wait();
                uint32_t w[2][2];
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int nt = 0; nt < 2; nt++) {
                        int v[4];
#pragma unroll
                        for (int c = 0; c < 4; c++) {
                            v[c] = tdiv1<FAST1>((int)r[mt][4 * nt + c] + off1[nt][c & 1], mg1[nt][c & 1], sh1[nt][c & 1]);
                            if (K::T % 32 != 0 && 32 * u + 16 * mt + g + 8 * (c >> 1) >= K::T) v[c] = 0;
                        }
                        w[mt][nt] = pack4_sat(v[0], v[1], v[2], v[3]);
                        if (TAPS && tap1) {
                            const long long b = first + (long long)n * stride;
#pragma unroll
                            for (int c = 0; c < 4; c++) {
                                const int f = 8 * nt + 2 * q + (c & 1), t = 32 * u + 16 * mt + g + 8 * (c >> 1);
                                if (t < K::T) tap1[((size_t)b * 16 + f) * K::T + t] = (int8_t)clip8(v[c]);
                            }
                        }
                    }
                stsm_t8_x4(yb + 16 * u, w[0][0], w[0][1], w[1][0], w[1][1]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(sbase + K::BAR_TEMPTY + 8 * ab);
                mbar_arrive(sbase + K::BAR_FULL_Y1 + 8 * buf);
            }
        }
      } else {
        // W1 as the B operand: [block][filter n-tile][lane] = (channels 4q..4q+3 of the block, filter 8 nt + g)
        uint32_t wfr[K::KB16][2];
#pragma unroll
        for (int kb = 0; kb < K::KB16; kb++) {
            wfr[kb][0] = net->w1blk[kb][0][lane];
            wfr[kb][1] = net->w1blk[kb][1][lane];
        }
        // accumulator columns 2q, 2q+1 of n-tile nt <-> filters 8 nt + 2q, + 1
        int off1[2][2], mg1[2][2], sh1[2][2];
#pragma unroll
        for (int nt = 0; nt < 2; nt++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                off1[nt][e] = net->off1[8 * nt + 2 * q + e];
                mg1[nt][e] = net->m1[8 * nt + 2 * q + e];
                sh1[nt][e] = net->sh1[8 * nt + 2 * q + e];
            }
        // REAL: the operand rows come from this warp's scratch ([channel][48 B], natural channel order)
        uint32_t xoff[K::KP + 1];
#pragma unroll
        for (int s = 0; s < K::KP; s++)
            xoff[s] = K::REAL ? (uint32_t)(K::OFF_SCR + warp * K::SCR_WARP + (16 * (lane >> 4) + (lane & 15)) * K::SCR_ROW)
                              : l1_row_offset<K>(2 * s + (lane >> 4), lane & 15);
        if (K::KR) xoff[K::KP] = K::REAL ? (uint32_t)(K::OFF_SCR + warp * K::SCR_WARP + (lane & 15) * K::SCR_ROW)
                                         : l1_row_offset<K>(K::KB16 - 1, lane & 15);
        static_assert(!K::REAL || K::KB16 <= 2, "scratch holds 32 channels");
        // stmatrix.trans: matrix j = lane >> 3 = 2 mt + nt holds filters 8 nt .. + 7, samples 16 mt .. + 15 of the
        // unit, i.e. chunk 2 (u + 1) + mt = plane mt, index u + 1
        const uint32_t st_off = ((lane & 7) + 8 * ((lane >> 3) & 1)) * K::Y1S + (lane >> 4) * K::PLANE + 16;

        for (int n = 0; n < nloc; n++) {
            const int st = n % K::S, buf = n % kNY1;
            PWAIT(0, sbase + K::BAR_FULL_X + 8 * st, (n / K::S) & 1);
            PWAIT(1, sbase + K::BAR_EMPTY_Y1 + 8 * buf, ((n / kNY1) & 1) ^ 1);
            const uint32_t xb = K::REAL ? sbase : sbase + K::OFF_XS + st * K::XS_BYTES;
            const uint32_t yb = sbase + K::OFF_Y1 + buf * K::Y1_BYTES + st_off;
            // REAL: byte misalignment of row `lane` of this trial (x itself is 16-byte aligned)
            const long long bq = first + (long long)n * stride;
            const int rsh = K::REAL ? (int)((bq * (long long)(K::TRIAL_BYTES & 15) + (long long)lane * (K::T & 15)) & 15) : 0;
#pragma unroll 1
            // the warp that gets one unit fewer rotates with the trial: A warps only meet at the y1 barriers
            for (int u = MIBMI_ROTATE ? (warp + n) % kAW : warp; u < K::UNITS; u += kAW) {
                if (K::REAL) {
                    // shift 32 samples of every channel into place: 9 aligned words -> 8 funnel shifts -> scratch
                    __syncwarp();
                    if (lane < K::C) {
                        const uint32_t* src = reinterpret_cast<const uint32_t*>(
                            smem + K::OFF_XS + st * K::XS_BYTES + lane * K::XP + 32 * u + (rsh & ~3));
                        uint32_t w[9], o[8];
#pragma unroll
                        for (int i = 0; i < 9; i++) w[i] = src[i];
                        const int bs = 8 * (rsh & 3);
#pragma unroll
                        for (int i = 0; i < 8; i++) o[i] = __funnelshift_r(w[i], w[i + 1], bs);
                        uint4* dst = reinterpret_cast<uint4*>(smem + K::OFF_SCR + warp * K::SCR_WARP + lane * K::SCR_ROW);
                        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
                        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
                    }
                    __syncwarp();
                }
                const uint32_t xu = K::REAL ? 0u : 32u * u;
                // D[16 samples x 8 filters] += X^T[16 samples x 32 ch] * W1^T[32 ch x 8 filters]: the four
                // registers of one transposing ldmatrix.x2 are exactly one A fragment
                int acc[2][2][4];
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int nt = 0; nt < 2; nt++) {
                        acc[mt][nt][0] = acc[mt][nt][2] = off1[nt][0];
                        acc[mt][nt][1] = acc[mt][nt][3] = off1[nt][1];
                    }
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
#pragma unroll
                    for (int s = 0; s < K::KP; s++) {
                        uint32_t d[4];
                        ldsm_t_x2(d, xb + xoff[s] + xu + 16 * mt);
                        mma_k32(acc[mt][0], d, wfr[2 * s][0], wfr[2 * s + 1][0]);
                        mma_k32(acc[mt][1], d, wfr[2 * s][1], wfr[2 * s + 1][1]);
                    }
                    if (K::KR) {
                        uint32_t d[2];
                        ldsm_t_x1(d, xb + xoff[K::KP] + xu + 16 * mt);
                        mma_k16(acc[mt][0], d[0], d[1], wfr[K::KB16 - 1][0]);
                        mma_k16(acc[mt][1], d[0], d[1], wfr[K::KB16 - 1][1]);
                    }
                }
                // requant; thread (g, q) owns samples 32u + 16mt + g, + 8 of filters 8nt + 2q, + 1
                uint32_t w[2][2];
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int nt = 0; nt < 2; nt++) {
                        int v[4];
#pragma unroll
                        for (int c = 0; c < 4; c++) {
                            v[c] = tdiv1<FAST1>(acc[mt][nt][c], mg1[nt][c & 1], sh1[nt][c & 1]);
                            // samples past T (only in the last unit) are layer 2's zero padding
                            if (K::T % 32 != 0 && 32 * u + 16 * mt + g + 8 * (c >> 1) >= K::T) v[c] = 0;
                        }
                        w[mt][nt] = pack4_sat(v[0], v[1], v[2], v[3]);
                        if (TAPS && tap1) {
                            const long long b = first + (long long)n * stride;
#pragma unroll
                            for (int c = 0; c < 4; c++) {
                                const int f = 8 * nt + 2 * q + (c & 1), t = 32 * u + 16 * mt + g + 8 * (c >> 1);
                                if (t < K::T) tap1[((size_t)b * 16 + f) * K::T + t] = (int8_t)clip8(v[c]);
                            }
                        }
                    }
                stsm_t8_x4(yb + 16 * u, w[0][0], w[0][1], w[1][0], w[1][1]);
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(sbase + K::BAR_EMPTY_X + 8 * st);
                mbar_arrive(sbase + K::BAR_FULL_Y1 + 8 * buf);
            }
        }
      }
    } else if (warp < kAW + kBW) {
        // =====================================================================================
        // B: layers 2 + 3 of channel k
        // =====================================================================================
        const int k = warp - kAW;
        uint32_t bt[4][3][2];
#pragma unroll
        for (int v = 0; v < 4; v++)
#pragma unroll
            for (int s = 0; s < 3; s++)
#pragma unroll
                for (int h = 0; h < 2; h++) bt[v][s][h] = net->w2frag[k][v * 6 + s * 2 + h][lane];
        const int thr = net->thr2[k], sig = net->sig2[k], soff = net->soff2[k], m2 = net->m2[k], sh2 = net->sh2[k];
        uint32_t b3[2][2];
#pragma unroll
        for (int v = 0; v < 2; v++)
#pragma unroll
            for (int h = 0; h < 2; h++) b3[v][h] = net->w3frag[k][v * 2 + h][lane];
        const int m3 = net->m3, sh3 = net->sh3;
        // ldmatrix.x4 lane addressing: matrix j = lane >> 3, row (lane & 7) + 8 (j & 1), k-half j >> 1
        const int lrow = (lane & 7) + 8 * ((lane >> 3) & 1);
        const uint32_t l2_off = k * K::Y1S + (lane >> 4) * K::PLANE + 16 * lrow;
        const int l3_row = lrow < K::NW3 ? lrow : K::NW3 - 1;
        const uint32_t l3_addr = sbase + K::OFF_Y2 + k * K::Y2S + 16 * (l3_row + (lane >> 4));
        uint8_t* y2row = smem + K::OFF_Y2 + k * K::Y2S;
        // layer-3 stores: sample t = 16 i + 8 v + 2 q (+1) goes to chunk 2 (i >> 1) + (q >> 1), position
        // 8 (q & 1) + 4 (i & 1) + 2 v: the order layer 4 wants (pooling group = MMA column pair)
        const uint32_t y3_off = k * K::Y3S + 16 * (q >> 1) + 8 * (q & 1);

        for (int n = 0; n < nloc; n++) {
            const int buf = n % kNY1, par = (n / kNY1) & 1;
            const int buf3 = n % kNY3, par3 = (n / kNY3) & 1;
            const long long b = first + (long long)n * stride;
            PWAIT(0, sbase + K::BAR_FULL_Y1 + 8 * buf, par);
            const uint32_t ya = sbase + K::OFF_Y1 + buf * K::Y1_BYTES + l2_off;
#pragma unroll 1
            for (int mt = 0; mt < K::MT EXP_SKIP_L2; mt++) {
                int acc[4][4];
#pragma unroll
                for (int v = 0; v < 4; v++) acc[v][0] = acc[v][1] = acc[v][2] = acc[v][3] = 0;
#pragma unroll
                for (int s = 0; s < 3; s++) {
                    uint32_t a[4];
                    ldsm_x4(a, ya + 16 * (16 * mt + s));
#pragma unroll
                    for (int v = 0; v < 4; v++) mma_k32(acc[v], a, bt[v][s][0], bt[v][s][1]);
                }
#pragma unroll
                for (int hf = 0; hf < 2; hf++) {
                    int s = 0;
#pragma unroll
                    for (int v = 0; v < 4; v++) s += max(acc[v][2 * hf], thr) + max(acc[v][2 * hf + 1], thr);
                    const int p = 4 * (16 * mt + g + 8 * hf) + q;     // pooled sample index
                    const int o = clip8(tdiv_magic(s * sig + soff, m2, sh2));
                    if (p < K::T8) {
                        y2row[8 + p] = (uint8_t)o;
                        if (TAPS && tap2) tap2[((size_t)b * 16 + k) * K::T8 + p] = (int8_t)o;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + K::BAR_EMPTY_Y1 + 8 * buf);
            PWAIT(1, sbase + K::BAR_EMPTY_Y3 + 8 * buf3, par3 ^ 1);
            // ---- layer 3
            if (EXP_DO_L3) {
                uint32_t a[4];
                ldsm_x4(a, l3_addr);
                int c3[2][4];
#pragma unroll
                for (int v = 0; v < 2; v++) {
                    c3[v][0] = c3[v][1] = c3[v][2] = c3[v][3] = 0;
                    mma_k32(c3[v], a, b3[v][0], b3[v][1]);
                }
                uint8_t* y3b = smem + K::OFF_Y3 + buf3 * K::Y3_BYTES + y3_off;
#pragma unroll
                for (int half = 0; half < (K::NW3 > 8 ? 2 : 1); half++) {
                    const int i = g + 8 * half;
#pragma unroll
                    for (int v = 0; v < 2; v++) {
                        const int t = 16 * i + 8 * v + 2 * q;
                        const int lo = tdiv_magic(c3[v][2 * half], m3, sh3), hi = tdiv_magic(c3[v][2 * half + 1], m3, sh3);
                        const uint32_t w = pack2_sat(lo, hi);
                        if (t < K::T8) {
                            *reinterpret_cast<uint16_t*>(y3b + 32 * (i >> 1) + 4 * (i & 1) + 2 * v) = (uint16_t)w;
                            if (TAPS && tap3) {
                                tap3[((size_t)b * 16 + k) * K::T8 + t] = (int8_t)(w & 0xff);
                                tap3[((size_t)b * 16 + k) * K::T8 + t + 1] = (int8_t)((w >> 8) & 0xff);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + K::BAR_FULL_Y3 + 8 * buf3);
        }
    } else if (warp == kAW + kBW + kCW) {
        // =====================================================================================
        // P: TMA producer -- refill the stage of trial n with trial n + S as soon as layer 1 released it
        // =====================================================================================
        auto issue = [&](int st, int n) {
            const size_t toff = (size_t)(first + (long long)n * stride) * K::TRIAL_BYTES;
            const uint32_t dst = sbase + K::OFF_XS + st * K::XS_BYTES;
            const uint32_t bar = sbase + K::BAR_FULL_X + 8 * st;
            if (K::REAL) {
                // lane c copies row c from the 16-byte boundary below it (x is 16-byte aligned).  No byte outside the batch
                // is read: the copy of the last row of the LAST trial stops at the 16-byte boundary below the end of the
                // batch and its lane fetches the remaining <= 15 bytes with byte loads
                const size_t off = toff + (size_t)lane * K::T, a0 = off & ~(size_t)15, end = (size_t)B * K::TRIAL_BYTES;
                size_t a1 = (off + K::T + 15) & ~(size_t)15;
                if (lane < K::C && a1 > end) {
                    a1 = end & ~(size_t)15;
                    for (size_t i = a1; i < end; i++) smem[K::OFF_XS + st * K::XS_BYTES + lane * K::XP + (i - a0)] = (uint8_t)x[i];
                }
                const uint32_t bytes = lane < K::C ? (uint32_t)(a1 - a0) : 0u;
                const uint32_t total = __reduce_add_sync(0xffffffffu, bytes);
                __syncwarp();                                      // the byte stores are ordered before lane 0's arrive (release)
                if (lane == 0) mbar_expect_tx(bar, total);
                __syncwarp();
                if (lane < K::C) bulk_g2s(dst + lane * K::XP, x + a0, bytes, bar);
            } else if (K::UMMA) {
                if (lane == 0) {
                    // 128-sample tiles [C][128] of the [B * C][T] tensor; columns past T are zero-filled
                    mbar_expect_tx(bar, K::TILES * K::C * 128);
#pragma unroll
                    for (int j = 0; j < K::TILES; j++)
                        tma_load_2d(dst + j * K::TILE_BYTES, &tmx, 128 * j, (int)((first + (long long)n * stride) * K::C), bar);
                }
            } else if (lane == 0) {
                mbar_expect_tx(bar, K::TRIAL_BYTES);
                bulk_g2s(dst, x + toff, K::CA * K::T, bar);
                if (K::CB > 0) bulk_g2s(dst + K::ROFF, x + toff + K::CA * K::T, K::CB * K::T, bar);
            }
        };
        if (K::REAL || lane == 0) {
            for (int s = 0; s < K::S && s < nloc; s++) issue(s, s);
            for (int n = 0; n + K::S < nloc; n++) {
                const int st = n % K::S;
                PWAIT(0, sbase + K::BAR_EMPTY_X + 8 * st, (n / K::S) & 1);
                issue(st, n + K::S);
            }
        }
    } else if (warp == kAW + kBW + kCW + 1) {
        // =====================================================================================
        // M: tcgen05.mma issuer (one thread).  Per trial: TILES x KSTEPS instructions
        //    D[128 samples x 16 filters] (+)= Xt[128 x 32 ch] * W1t[32 x 16], then two commits: the input stage
        //    is free (no A-warp touches it any more) and the accumulators are complete
        // =====================================================================================
        if (K::UMMA && lane == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(128, 16);
            for (int n = 0; n < nloc; n++) {
                const int st = n % K::S, ab = n & 1;
                PWAIT(0, sbase + K::BAR_FULL_X + 8 * st, (n / K::S) & 1);
                PWAIT(1, sbase + K::BAR_TEMPTY + 8 * ab, ((n >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t xs = sbase + K::OFF_XS + st * K::XS_BYTES;
#pragma unroll
                for (int j = 0; j < K::TILES; j++)
#pragma unroll
                    for (int ks = 0; ks < K::KSTEPS; ks++)
                        umma_i8(tmem_base + ab * K::TMEM_COLS_TRIAL + 16 * j,
                                umma_desc(xs + j * K::TILE_BYTES + ks * 4096, 16, 1024, 2),      // MN-major, 128B swizzle
                                umma_desc(sbase + K::OFF_W1 + ks * 256, 128, K::W1_SBO, 0),      // K-major core matrices
                                idesc, ks > 0);
                umma_commit(sbase + K::BAR_EMPTY_X + 8 * st);
                umma_commit(sbase + K::BAR_TFULL + 8 * ab);
            }
        }
    } else {
        // =====================================================================================
        // C: layers 4 + 5; warp j handles the trials with n = j (mod 2), i.e. y3 buffer j
        // =====================================================================================
        const int cj = warp - (kAW + kBW);
        const uint32_t w4a0 = net->w4frag[0][lane], w4a1 = net->w4frag[1][lane];
        const int thr_lo = net->thr4[g], thr_hi = net->thr4[g + 8];
        const int sig_lo = net->sig4[g], sig_hi = net->sig4[g + 8];
        const int soff_lo = net->soff4[g], soff_hi = net->soff4[g + 8];
        const int m4_lo = net->m4[g], m4_hi = net->m4[g + 8], sh4_lo = net->sh4[g], sh4_hi = net->sh4[g + 8];
        const int b5 = net->b5[g & 3], m5 = net->m5, sh5 = net->sh5;
        const uint32_t l4_off = (lane & 15) * K::Y3S + 16 * (lane >> 4);
        uint8_t* y4 = smem + K::OFF_Y4 + cj * K::Y4_BYTES;
        uint32_t w5a[K::KST5 <= 4 ? K::KST5 : 1][2];
        if (K::KST5 <= 4) {
#pragma unroll
            for (int ks = 0; ks < K::KST5; ks++) { w5a[ks][0] = net->w5frag[ks][0][lane]; w5a[ks][1] = net->w5frag[ks][1][lane]; }
        }
        const uint32_t* y4w = reinterpret_cast<const uint32_t*>(y4);

        for (int n = cj; n < nloc; n += kCW) {
            const int buf = n % kNY3, par = (n / kNY3) & 1;
            const long long b = first + (long long)n * stride;
            PWAIT(0, sbase + K::BAR_FULL_Y3 + 8 * buf, par);
            const uint32_t y3a = sbase + K::OFF_Y3 + buf * K::Y3_BYTES + l4_off;
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
                int s_lo = 0, s_hi = 0;
#pragma unroll
                for (int v = 0; v < 4; v++) {
                    s_lo += max(c[v][0], thr_lo) + max(c[v][1], thr_lo);
                    s_hi += max(c[v][2], thr_hi) + max(c[v][3], thr_hi);
                }
                const int grp = 4 * s4 + q;
                const int o_lo = clip8(tdiv_magic(s_lo * sig_lo + soff_lo, m4_lo, sh4_lo));
                const int o_hi = clip8(tdiv_magic(s_hi * sig_hi + soff_hi, m4_hi, sh4_hi));
                if (grp < K::T64) {
                    y4[g * K::T64P + grp] = (uint8_t)o_lo;
                    y4[(g + 8) * K::T64P + grp] = (uint8_t)o_hi;
                    if (TAPS && tap4) {
                        tap4[((size_t)b * 16 + g) * K::T64 + grp] = (int8_t)o_lo;
                        tap4[((size_t)b * 16 + g + 8) * K::T64 + grp] = (int8_t)o_hi;
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + K::BAR_EMPTY_Y3 + 8 * buf);
            // ---- layer 5: D[class][0] = W5[class][k, t] . y4[k, t]; only MMA column 0 carries data
            int z[4] = {0, 0, 0, 0};
#pragma unroll
            for (int ks = 0; ks < K::KST5; ks++) {
                const uint32_t a[4] = {K::KST5 <= 4 ? w5a[K::KST5 <= 4 ? ks : 0][0] : __ldg(&net->w5frag[ks][0][lane]), 0u,
                                       K::KST5 <= 4 ? w5a[K::KST5 <= 4 ? ks : 0][1] : __ldg(&net->w5frag[ks][1][lane]), 0u};
                const uint32_t b0 = g == 0 ? y4w[8 * ks + q] : 0u, b1 = g == 0 ? y4w[8 * ks + 4 + q] : 0u;
                mma_k32(z, a, b0, b1);
            }
            if (q == 0 && g < 4) {
                const int8_t o = (int8_t)clip8(tdiv_magic(z[0] + b5, m5, sh5));
                logits[(size_t)b * 4 + g] = o;
                // fused all-gather: the same 4 bytes go to every peer's buffer (NVLink P2P stores)
                for (int i = 0; i < peers.n; i++) peers.p[i][(size_t)b * 4 + g] = o;
            }
            __syncwarp();   // y4 is rewritten by the next trial
        }
    }
    PROF_END
    if (K::UMMA) {
        tc_fence_before();
        __syncthreads();
        if (warp == kAW + kBW + kCW + 1) tmem_dealloc(tmem_base, K::TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------
tmap_encode_t tmap_encoder() {
    static tmap_encode_t fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess ||
            qr != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (tmap_encode_t)p;
    }();
    return fn;
}

template <int C, int T>
static cudaError_t launch_pipe_instance(const PipeNet* net_dev, bool fast1, const int8_t* x, long long B,
                                        int8_t* logits, const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4,
                                        int num_sms, cudaStream_t stream) {
    // deepest input ring (<= MIBMI_STAGES) that fits the 227 KB of shared memory
    constexpr int S = PCfg<C, T, MIBMI_STAGES>::SMEM_RAW <= 227 * 1024 ? MIBMI_STAGES
                    : PCfg<C, T, MIBMI_STAGES - 1>::SMEM_RAW <= 227 * 1024 ? MIBMI_STAGES - 1
                    : PCfg<C, T, MIBMI_STAGES - 2>::SMEM_RAW <= 227 * 1024 ? MIBMI_STAGES - 2 : 3;
    using K = PCfg<C, T, S>;
    CUtensorMap tm;
    memset(&tm, 0, sizeof(tm));
    if (K::UMMA) {
        tmap_encode_t enc = tmap_encoder();
        if (!enc) return cudaErrorNotSupported;
        cuuint64_t dims[2] = {(cuuint64_t)T, (cuuint64_t)(B * C)};
        cuuint64_t strides[1] = {(cuuint64_t)T};
        cuuint32_t box[2] = {128, (cuuint32_t)C};
        cuuint32_t estr[2] = {1, 1};
        if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<int8_t*>(x), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return cudaErrorInvalidValue;
    }
    const bool taps = t1 || t2 || t3 || t4;
    auto kern = taps ? pipe_forward_kernel<K, false, true>
                     : (fast1 ? pipe_forward_kernel<K, true, false> : pipe_forward_kernel<K, false, false>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
    if (e != cudaSuccess) return e;
    long long grid = num_sms;
    if (grid > B) grid = B;
    kern<<<(unsigned)grid, kPipeThreads, K::SMEM, stream>>>(tm, net_dev, x, B, logits, peers, t1, t2, t3, t4);
    return cudaGetLastError();
}

#define MIBMI_PIPE_SHAPES(X) X(64, 480) X(38, 480) X(24, 480) X(19, 480) X(16, 480) X(8, 480) X(22, 1125)

bool pipe_has_instance(int C, int T) {
#define X(c, t) if (C == c && T == t) return true;
    MIBMI_PIPE_SHAPES(X)
#undef X
    return false;
}

cudaError_t launch_pipe(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B,
                        int8_t* logits, const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                        cudaStream_t stream) {
#define X(c, t) if (C == c && T == t) return launch_pipe_instance<c, t>(net_dev, fast1, x, B, logits, peers, t1, t2, t3, t4, num_sms, stream);
    MIBMI_PIPE_SHAPES(X)
#undef X
    return cudaErrorInvalidValue;
}

}  // namespace mibmi

#ifdef MIBMI_PROF
extern "C" __attribute__((visibility("default"))) int mibmi_debug_prof(long long* out_host) {
    return (int)cudaMemcpyFromSymbol(out_host, mibmi::g_prof, sizeof(mibmi::g_prof));
}
#endif
