// Tensor-core-resident int8 edgeEEGNet forward for sm_100a: layers 1 AND 2 run on tcgen05.mma (kind::i8) with
// their int32 accumulators in TMEM; the warps only requantise.  (Successor of mibmi_pipe.cu for shapes whose
// rows are 16-byte aligned; measured motivation in DESIGN.md 4.2: the mma.sync kernel is bound by the dependent
// latency of its layer-2 warps, not by any pipe.)
//
// One persistent CTA per SM (23 warps) works on QUADS of 4 trials:
//   P   warp 20   2-D TMA tiles [C][128 samples], 128-byte swizzle, 4 per trial, 3-stage ring
//   M1  warp 21   layer 1: per tile D[128 samples x 16 filters] = Xt[128 x C] * W1t (A = the swizzled TMA tile read
//                 MN-major, B = W1 K-major core matrices); accumulators of two trials in TMEM columns 0..127
//                 -- edge-eegnet_wolf/src/cl/net/layer1.c:87-96
//   A   warps 0-7 layer-1 epilogue: tcgen05.ld.16x256b (= the m16n8 fragment), + BN offset, magic division,
//                 saturating pack, stmatrix.trans into y1[filter][32 zeros | trial | 32 zeros | trial ...]
//   M2  warp 22   layer 2 (depthwise 64-tap conv) of a whole quad per temporal channel as 3 instructions:
//                 D[128 rows x 16 outputs] = Hankel(y1 row)[128 x 96] * Toeplitz(taps)[96 x 16], where the Hankel
//                 operand is the y1 byte array itself described with overlapping strides (row m = bytes 16 m ..,
//                 LBO = 16 B, SBO = 128 B): no im2col.  Rows 32 s + i of the tile = outputs 16 i .. 16 i + 15 of the
//                 trial in slot s; the 32 zero bytes between trials are both trials' padding
//                 (gen_net_header.py:145-146).  TMEM columns 128 + 16 k -- layer2.c:87-118 (REORDER_BN)
//   B   warps 8-15 layer-2 epilogue (thread = TMEM lane = 16 consecutive conv outputs = two pooling groups:
//                 threshold, sum, requant, no shuffles), then layer 3 with mma.sync where the 16 MMA rows are the
//                 4 trials x 4 blocks of 16 samples -- layer3.c:70
//   C   warps 16-19 layers 4 + 5 of one trial each, as in mibmi_pipe.cu -- layer4.c:99-140, layer5.c:81-89
// Hand-off is mbarrier-only; tcgen05.commit frees input stages / y1 buffers and publishes accumulators.
// Bit-identical to the reference (tests/test_parity_gpu.py).
#include <cuda.h>
#include <string.h>

#include "mibmi_common.cuh"
#include "mibmi_kernels.h"
#include "mibmi_ptx.cuh"
#include "mibmi_tmap.h"

namespace mibmi {

#ifndef MIBMI_TC_SLEEP_NS
#define MIBMI_TC_SLEEP_NS 64
#endif
#ifndef MIBMI_TC_WAIT
#define MIBMI_TC_WAIT 0                   // mbar_wait_mode: 0 = test_wait + nanosleep, 1 = test_wait then try_wait, 2 = try_wait
#endif
#ifdef MIBMI_PROF
__device__ long long g_prof_tc[148 * 32 * 4];
#define TWAIT(slot, ...) do { if (!mbar_test(__VA_ARGS__)) { const long long t0_ = clock64(); mbar_wait_mode<MIBMI_TC_WAIT>(__VA_ARGS__, MIBMI_TC_SLEEP_NS); pw_[slot] += clock64() - t0_; } } while (0)
#define TPROF_BEGIN long long pw_[3] = {0, 0, 0}; const long long pt0_ = clock64();
#define TPROF_END if ((threadIdx.x & 31) == 0) { long long* o_ = g_prof_tc + (blockIdx.x * 32 + (threadIdx.x >> 5)) * 4; \
    o_[0] = clock64() - pt0_; o_[1] = pw_[0]; o_[2] = pw_[1]; o_[3] = pw_[2]; }
#else
#define TWAIT(slot, ...) mbar_wait_mode<MIBMI_TC_WAIT>(__VA_ARGS__, MIBMI_TC_SLEEP_NS)
#define TPROF_BEGIN
#define TPROF_END
#endif

#ifndef MIBMI_TC_AW
#define MIBMI_TC_AW 8                     // layer-1 epilogue warps: a multiple of 4 (one per TMEM lane quarter)
#endif
constexpr int kTcAW = MIBMI_TC_AW, kTcBW = 8, kTcCW = 4, kTcNA = kTcAW / 4;
static_assert(kTcAW % 4 == 0 && kTcAW >= 4, "layer-1 epilogue warps come in sets of four TMEM lane quarters");
constexpr int kTcWarpP = kTcAW + kTcBW + kTcCW, kTcWarpM1 = kTcWarpP + 1, kTcWarpM2 = kTcWarpP + 2;
constexpr int kTcThreads = (kTcWarpM2 + 1) * 32;

template <int C_, int T_>
struct TCfg {
    static constexpr int C = C_, T = T_, SX = 3;
    static constexpr int TILES = (T + 127) / 128, KPAD = (C + 31) / 32 * 32, KSTEPS = KPAD / 32;
    static constexpr int TILE_BYTES = KPAD * 128, XS_BYTES = TILES * TILE_BYTES;
    static constexpr int T8 = T / 8, T64 = T8 / 8, ROWS = T / 16;          // layer-2 rows (16 outputs) per trial
    static constexpr int SLOT = 512;                                        // trial pitch in a y1 row: 32 MMA rows
    static constexpr int CHS = 32 + 8 * SLOT + 16;                          // y1 row of one filter: 2 quads; / 16 odd
    static constexpr int NW3 = (T8 + 15) / 16;                              // layer-3 blocks of 16 samples
    static constexpr int Y2S = 80, Y2_BUF = 4 * 16 * Y2S;
    static constexpr int Y3S = 80, Y3_TRIAL = 16 * Y3S;
    static constexpr int NST = (T64 + 3) / 4, T64P = (T64 + 1) / 2 * 2, KST5 = T64P / 2;
    static constexpr int Y4_BYTES = ((16 * T64P + 15) / 16) * 16;
    static constexpr int W2_CH = 16 * 96;                                   // Toeplitz B operand of one channel
    // Wolf order with rows that are not a multiple of 16 bytes (WOLF = 3): super-rows of WS samples (PipeNet::w1x)
    static constexpr int CA = (C + 3) / 4 * 4;
    static constexpr int WS = CA % 16 == 0 ? 1 : (CA % 8 == 0 ? 2 : 4);    // 16 / gcd(CA, 16)
    static constexpr int RB3 = WS * CA, NCH3 = RB3 / 16, KS3 = (RB3 + 31) / 32, KP3 = 32 * KS3;
    static constexpr int TILE3 = 2 * KS3 * 2048;                            // [chunk][128 super-rows][16 B], padded to whole k-steps
    static constexpr int TILES3 = (T / WS + 127) / 128;
    // TMEM columns
    static constexpr int TM_L1 = 0, TM_L1_TRIAL = 16 * TILES, TM_L2 = 128;
    // barriers
    static constexpr int BAR_FULL_X = 0, BAR_EMPTY_X = 8 * SX, BAR_T1_FULL = 16 * SX, BAR_T1_EMPTY = BAR_T1_FULL + 16,
                         BAR_Y1_FULL = BAR_T1_EMPTY + 16, BAR_Y1_EMPTY = BAR_Y1_FULL + 16, BAR_T2_FULL = BAR_Y1_EMPTY + 16,
                         BAR_T2_EMPTY = BAR_T2_FULL + 16, BAR_Y2_FULL = BAR_T2_EMPTY + 16, BAR_Y3_FULL = BAR_Y2_FULL + 16,
                         BAR_Y3_EMPTY = BAR_Y3_FULL + 16, BAR_END = BAR_Y3_EMPTY + 16;
    // shared memory map
    static constexpr int OFF_TMEM = 248;
    static constexpr int OFF_C2 = 256;                                      // int [16][8]: thr, sig, soff, m, sh
    static constexpr int OFF_W1 = 1024;
    static constexpr int OFF_W2 = 2048;
    static constexpr int OFF_W1X = OFF_W2 + 16 * W2_CH;                     // block-structured W1 of the super-row form (6 KB)
    static constexpr int OFF_XS = ((OFF_W1X + kW1X + 1023) / 1024) * 1024;
    static constexpr int OFF_Y1 = OFF_XS + SX * XS_BYTES;
    static constexpr int OFF_Y2 = OFF_Y1 + 16 * CHS;
    static constexpr int OFF_Y3 = OFF_Y2 + 2 * Y2_BUF;
    static constexpr int OFF_Y4 = OFF_Y3 + 8 * Y3_TRIAL;
    static constexpr int OFF_LG = ((OFF_Y4 + kTcCW * Y4_BYTES + 15) / 16) * 16;        // 2 x 16 B: the logits of a quad, for the peer stores
    static constexpr int SMEM = ((OFF_LG + 32 + 127) / 128) * 128;
    static_assert(T % 32 == 0 && T <= SLOT - 32, "trial fits a 512-byte slot with 32 zero bytes");
    static_assert(TM_L1 + 2 * TM_L1_TRIAL <= TM_L2, "TMEM map");
    static_assert(NW3 <= 4, "layer 3: 4 trials x 4 blocks fill the 16 MMA rows");
    static_assert(KST5 <= 4, "classifier fragments live in registers");
    static_assert(BAR_END <= OFF_TMEM && 16 * KPAD <= 1024, "barrier / W1 area");
    static_assert(SMEM <= 227 * 1024, "shared memory");
    static_assert(WS == 1 || (RB3 % 16 == 0 && T % WS == 0 && 16 * WS * KP3 <= kW1X && TILES3 * TILE3 <= XS_BYTES &&
                              TILES3 * 16 * WS <= TM_L1_TRIAL && OFF_W1X % 128 == 0), "super-row form");
};

template <bool FAST>
__device__ __forceinline__ int tc_div1(int x, int m, int s) {
    if (FAST) return __mulhi(x, m) + (int)((unsigned)x >> 31);
    return (__mulhi(x, m) >> s) + (int)((unsigned)x >> 31);
}

// layer 1 with the BN offset folded into the multiplication (FAST form only): hi32((x + off) m) = hi32(x m + off m) -- one
// multiply-add with a 64-bit constant instead of an add and a multiply-high; the quotient's sign bit stands in for (x + off < 0)
// (m > 0, so hi32 is negative exactly when x + off is).  |x + off| < 2^29 and m < 2^31 keep the 64-bit sum exact.
#ifndef MIBMI_TC_FOLD
#define MIBMI_TC_FOLD 0   // measured r1g: no gain at 64 x 480 (183 M trials/s either way) and 68 bytes of spills at 80 registers
#endif
__device__ __forceinline__ int tc_div1_fold(int x, int m, long long offm) {
    const int hi = (int)(((long long)x * m + offm) >> 32);
    return hi + (int)((unsigned)hi >> 31);
}

__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t (&r)[16], uint32_t taddr) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr));
}
// instruction descriptor with both operands K-major (layer 2)
__device__ __forceinline__ constexpr uint32_t umma_idesc_i8_kk(int M, int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// WOLF: x is [B][T][C_ALIGN] (the order of net_model_compute, gen_input_header.py:75-76), already K-major for layer 1: a 3-D
// tensor map (16 bytes, time rows, 16-byte channel chunks) lands every tile as [chunk][128 rows][16 B] = K-major 8 x 16-byte core
// matrices (rows 16 B apart, SBO 128 B, LBO 2048 B), no swizzle -- the descriptor form of the layer-2 Hankel operand (WOLF = 1).
// WOLF = 2 (rows of exactly 64 bytes): 2-D tiles [128 rows][64 B] with the 64-byte swizzle, K-major SWIZZLE_64B descriptor
// (SBO 512 B = 8 rows, the second k-step starts 32 B into the row): full-row TMA requests instead of 16-byte ones.
// PLAIN: the REORDER_BN-off variant of the pooled layers (layer2.c:139-210, layer4.c:99-140 -- the C, which does not clip the layer-4
// quotients): every accumulator is divided, u = tdiv(sig v + soff, |fac >> 3|), then ReLU (layer 2: clip to 127 first), sum of 8, >> 3.
template <class K, bool FAST1, bool TAPS, int WOLF, bool PLAIN = false>
__global__ void __launch_bounds__(kTcThreads, 1)
tc_forward_kernel(const __grid_constant__ CUtensorMap tmx, const PipeNet* __restrict__ net, long long B,
                  int8_t* __restrict__ logits, const PeerOut peers, int8_t* __restrict__ tap1, int8_t* __restrict__ tap2,
                  int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const uint32_t sbase = smem_u32(smem);
    const long long nquads_total = (B + 3) / 4;
    const long long firstq = blockIdx.x, strideq = gridDim.x;
    const int nq = firstq < nquads_total ? (int)((nquads_total - firstq + strideq - 1) / strideq) : 0;   // quads of this CTA
    const int ntr = 4 * nq;                                                                               // trial slots

    // ---- prologue ------------------------------------------------------------------------------------
    if (tid == 0) {
        for (int s = 0; s < K::SX; s++) {
            mbar_init(sbase + K::BAR_FULL_X + 8 * s, 1);
            mbar_init(sbase + K::BAR_EMPTY_X + 8 * s, 1);            // tcgen05.commit of M1
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(sbase + K::BAR_T1_FULL + 8 * i, 1);            // commit of M1
            mbar_init(sbase + K::BAR_T1_EMPTY + 8 * i, kTcAW);
            mbar_init(sbase + K::BAR_Y1_FULL + 8 * i, 4 * kTcAW);    // 4 trials x 8 layer-1 warps
            mbar_init(sbase + K::BAR_Y1_EMPTY + 8 * i, 1);           // commit of M2
            mbar_init(sbase + K::BAR_T2_FULL + 8 * i, 1);            // commit of M2 (per half of the channels)
            mbar_init(sbase + K::BAR_T2_EMPTY + 8 * i, kTcBW / 2);
            mbar_init(sbase + K::BAR_Y2_FULL + 8 * i, kTcBW);
            mbar_init(sbase + K::BAR_Y3_EMPTY + 8 * i, kTcCW);       // "y2 buffer consumed by layer 3"
        }
        mbar_fence_init();
    }
    for (int i = tid; i < 16 * K::KPAD / 4; i += kTcThreads)
        reinterpret_cast<uint32_t*>(smem + K::OFF_W1)[i] = reinterpret_cast<const uint32_t*>(net->w1umma)[i];
    for (int i = tid; i < 16 * K::W2_CH / 16; i += kTcThreads)
        reinterpret_cast<uint4*>(smem + K::OFF_W2)[i] = reinterpret_cast<const uint4*>(net->w2umma)[i];
    if (WOLF == 3)
        for (int i = tid; i < 16 * K::WS * K::KP3 / 16; i += kTcThreads)
            reinterpret_cast<uint4*>(smem + K::OFF_W1X)[i] = reinterpret_cast<const uint4*>(net->w1x)[i];
    for (int i = tid; i < 16 * 8; i += kTcThreads) {
        const int k = i >> 3, f = i & 7;
        reinterpret_cast<int*>(smem + K::OFF_C2)[i] = f == 0 ? net->thr2[k] : f == 1 ? net->sig2[k] : f == 2 ? net->soff2[k]
                                                      : f == 3 ? net->m2[k] : f == 4 ? net->sh2[k] : 0;
    }
    {
        uint4* z = reinterpret_cast<uint4*>(smem + K::OFF_Y1);     // zero gaps of y1, y2 / y3 padding
        for (int i = tid; i < (K::SMEM - K::OFF_Y1) / 16; i += kTcThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (warp == kTcWarpM1) tmem_alloc(sbase + K::OFF_TMEM, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<const uint32_t*>(smem + K::OFF_TMEM);
    TPROF_BEGIN

    if (warp < kTcAW) {
        // =====================================================================================
        // A: layer-1 epilogue.  TMEM lanes 32 (warp & 3) .. + 31 of tiles jb, jb + 2
        // =====================================================================================
        int off1[2][2], mg1[2][2], sh1[2][2];
#pragma unroll
        for (int nt = 0; nt < 2; nt++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                off1[nt][e] = net->off1[8 * nt + 2 * q + e];
                mg1[nt][e] = net->m1[8 * nt + 2 * q + e];
                sh1[nt][e] = net->sh1[8 * nt + 2 * q + e];
            }
        long long offm[2][2];
#pragma unroll
        for (int nt = 0; nt < 2; nt++)
#pragma unroll
            for (int e = 0; e < 2; e++) offm[nt][e] = (long long)off1[nt][e] * mg1[nt][e];
        // stmatrix.trans: matrix j = lane >> 3 = 2 mt + nt: filters 8 nt .. + 7 (rows), samples 16 mt .. + 15
        const uint32_t st_off = sbase + K::OFF_Y1 + ((lane & 7) + 8 * ((lane >> 3) & 1)) * K::CHS + 16 * (lane >> 4) + 32;
        const int qi = warp & 3;
        for (int n = 0; n < ntr; n++) {
            const int gq = n >> 2, s = n & 3, buf = gq & 1, ab = n & 1;
            TWAIT(0, sbase + K::BAR_T1_FULL + 8 * ab, (n >> 1) & 1);
            tc_fence_after();
            if (s == 0) TWAIT(1, sbase + K::BAR_Y1_EMPTY + 8 * buf, ((gq >> 1) & 1) ^ 1);
            const uint32_t yb = st_off + K::SLOT * (4 * buf + s);
            if (WOLF == 3) {
                // super-row form: TMEM lane = super-row of WS samples, column 16 p + f = (sample p of the row, filter f).  The warp takes
                // the quarter qi of tile jb3 = 32 super-rows = 32 WS samples; per half of 16 lanes the thread holds, for its four
                // filters and two rows, the WS adjacent samples of a row: ONE store of WS bytes each into y1[f][WS row ..]
                // tiles jb3 with (jb3 + n) % kTcNA == warp >> 2 (at most one per warp and trial: TILES3 <= 2 <= kTcNA)
                const int jb3 = (((warp >> 2) - n) % kTcNA + kTcNA) % kTcNA;
                const int sr0 = 128 * jb3 + 32 * qi;
                static_assert(K::WS == 1 || K::TILES3 <= kTcNA, "one super-row tile per warp and trial");
                if (jb3 < K::TILES3 && K::WS * sr0 < K::T) {
                    const uint32_t ta = tmem_base + ((uint32_t)(32 * qi) << 16) + K::TM_L1 + ab * K::TM_L1_TRIAL + 16 * K::WS * jb3;
                    const uint32_t y1s = sbase + K::OFF_Y1 + 32 + K::SLOT * (4 * buf + s);
#pragma unroll
                    for (int mt = 0; mt < 2; mt++) {
                        if (K::WS * (sr0 + 16 * mt) >= K::T) break;                    // warp-uniform
                        uint32_t r[K::WS][8];
#pragma unroll
                        for (int pp = 0; pp < K::WS; pp++) tmem_ld_16x256b_x2(r[pp], ta + 16 * pp + ((uint32_t)(16 * mt) << 16));
                        tmem_ld_wait();
#pragma unroll
                        for (int nt = 0; nt < 2; nt++)
#pragma unroll
                            for (int c = 0; c < 4; c++) {
                                const int f = 8 * nt + 2 * q + (c & 1), sr = sr0 + 16 * mt + g + 8 * (c >> 1);
                                int v[K::WS];
#pragma unroll
                                for (int pp = 0; pp < K::WS; pp++)
                                    v[pp] = (FAST1 && MIBMI_TC_FOLD) ? tc_div1_fold((int)r[pp][4 * nt + c], mg1[nt][c & 1], offm[nt][c & 1])
                                                                    : tc_div1<FAST1>((int)r[pp][4 * nt + c] + off1[nt][c & 1], mg1[nt][c & 1], sh1[nt][c & 1]);
                                if (K::WS * sr < K::T) {
                                    const uint32_t ya = y1s + f * K::CHS + K::WS * sr;
                                    if (K::WS == 2)
                                        asm volatile("st.shared.u16 [%0], %1;" ::"r"(ya), "h"((uint16_t)pack2_sat(v[0], v[K::WS > 1 ? 1 : 0])) : "memory");
                                    else
                                        asm volatile("st.shared.u32 [%0], %1;" ::"r"(ya), "r"(pack4_sat(v[0], v[K::WS > 1 ? 1 : 0], v[K::WS > 2 ? 2 : 0], v[K::WS > 3 ? 3 : 0])) : "memory");
                                    if (TAPS && tap1) {
                                        const long long b = 4 * (firstq + (long long)gq * strideq) + s;
                                        if (b < B)
#pragma unroll
                                            for (int pp = 0; pp < K::WS; pp++)
                                                tap1[((size_t)b * 16 + f) * K::T + K::WS * sr + pp] = (int8_t)clip8(v[pp]);
                                    }
                                }
                            }
                    }
                }
            } else {
            // the warp's (up to) two units of a pass are loaded from TMEM together and requantised interleaved: the
                // TMEM latency and the dependent division chains of one unit hide behind the other
    #pragma unroll 1
                for (int jb = (((warp >> 2) - n) % kTcNA + kTcNA) % kTcNA; jb < K::TILES; jb += 2 * kTcNA) {   // tiles jb, jb + kTcNA of a pass
                    const int u0 = 4 * jb + qi, u1 = u0 + 4 * kTcNA;
                    if (32 * u0 >= K::T) break;
                    const bool two = jb + kTcNA < K::TILES && 32 * u1 < K::T;
                    uint32_t r[2][2][8];
                    const uint32_t ta = tmem_base + ((uint32_t)(32 * qi) << 16) + K::TM_L1 + ab * K::TM_L1_TRIAL + 16 * jb;
                    tmem_ld_16x256b_x2(r[0][0], ta);
                    tmem_ld_16x256b_x2(r[0][1], ta + (16u << 16));
                    if (two) {
                        tmem_ld_16x256b_x2(r[1][0], ta + 16 * kTcNA);
                        tmem_ld_16x256b_x2(r[1][1], ta + 16 * kTcNA + (16u << 16));
                    }
                    tmem_ld_wait();
    #pragma unroll
                    for (int uu = 0; uu < 2; uu++) {
                        if (uu == 1 && !two) break;
                        const int u = uu ? u1 : u0;
                        uint32_t w[2][2];
    #pragma unroll
                        for (int mt = 0; mt < 2; mt++)
    #pragma unroll
                            for (int nt = 0; nt < 2; nt++) {
                                int v[4];
    #pragma unroll
                                for (int c = 0; c < 4; c++)
                                    v[c] = (FAST1 && MIBMI_TC_FOLD)
                                               ? tc_div1_fold((int)r[uu][mt][4 * nt + c], mg1[nt][c & 1], offm[nt][c & 1])
                                               : tc_div1<FAST1>((int)r[uu][mt][4 * nt + c] + off1[nt][c & 1], mg1[nt][c & 1], sh1[nt][c & 1]);
                                w[mt][nt] = pack4_sat(v[0], v[1], v[2], v[3]);
                                if (TAPS && tap1) {
                                    const long long b = 4 * (firstq + (long long)gq * strideq) + s;
                                    if (b < B)
    #pragma unroll
                                        for (int c = 0; c < 4; c++) {
                                            const int f = 8 * nt + 2 * q + (c & 1), t = 32 * u + 16 * mt + g + 8 * (c >> 1);
                                            tap1[((size_t)b * 16 + f) * K::T + t] = (int8_t)clip8(v[c]);
                                        }
                                }
                            }
                        stsm_t8_x4(yb + 32 * u, w[0][0], w[0][1], w[1][0], w[1][1]);
                    }
                }
            }
            tc_fence_before();
            fence_proxy_async();          // y1 is read by the tensor core (async proxy)
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(sbase + K::BAR_T1_EMPTY + 8 * ab);
                mbar_arrive(sbase + K::BAR_Y1_FULL + 8 * buf);
            }
        }
    } else if (warp == kTcWarpP) {
        // =====================================================================================
        // P: TMA producer
        // =====================================================================================
        if (lane == 0) {
            auto issue = [&](int st, int n) {
                const long long b = 4 * (firstq + (long long)(n >> 2) * strideq) + (n & 3);   // b >= B: zero filled
                const uint32_t dst = sbase + K::OFF_XS + st * K::XS_BYTES;
                const uint32_t bar = sbase + K::BAR_FULL_X + 8 * st;
                if (WOLF == 3) {
                    mbar_expect_tx(bar, K::TILES3 * K::NCH3 * 2048);
#pragma unroll
                    for (int j = 0; j < K::TILES3; j++)
                        tma_load_3d(dst + j * K::TILE3, &tmx, 0, (int)(b * (K::T / K::WS)) + 128 * j, 0, bar);
                } else if (WOLF == 1) {
                    mbar_expect_tx(bar, K::TILES * K::TILE_BYTES);
#pragma unroll
                    for (int j = 0; j < K::TILES; j++) tma_load_3d(dst + j * K::TILE_BYTES, &tmx, 0, (int)(b * K::T) + 128 * j, 0, bar);
                } else if (WOLF == 2) {
                    mbar_expect_tx(bar, K::TILES * K::TILE_BYTES);
#pragma unroll
                    for (int j = 0; j < K::TILES; j++) tma_load_2d(dst + j * K::TILE_BYTES, &tmx, 0, (int)(b * K::T) + 128 * j, bar);
                } else {
                    mbar_expect_tx(bar, K::TILES * K::C * 128);
#pragma unroll
                    for (int j = 0; j < K::TILES; j++) tma_load_2d(dst + j * K::TILE_BYTES, &tmx, 128 * j, (int)(b * K::C), bar);
                }
            };
            for (int s = 0; s < K::SX && s < ntr; s++) issue(s, s);
            for (int n = 0; n + K::SX < ntr; n++) {
                const int st = n % K::SX;
                TWAIT(0, sbase + K::BAR_EMPTY_X + 8 * st, (n / K::SX) & 1);
                issue(st, n + K::SX);
            }
        }
    } else if (warp == kTcWarpM1) {
        // =====================================================================================
        // M1: layer-1 MMAs, TILES x KSTEPS per trial
        // =====================================================================================
        {
            constexpr uint32_t idesc = WOLF == 3 ? umma_idesc_i8_kk(128, 16 * K::WS) : WOLF ? umma_idesc_i8_kk(128, 16) : umma_idesc_i8(128, 16);
            // descriptors: only the start-address field (low 14 bits, units of 16 B) changes between instructions
            const uint64_t da0 = (WOLF == 1 || WOLF == 3) ? umma_desc(sbase + K::OFF_XS, 2048, 128, 0)
                                 : WOLF == 2 ? umma_desc(sbase + K::OFF_XS, 16, 512, 4) : umma_desc(sbase + K::OFF_XS, 16, 1024, 2),
                           db0 = WOLF == 3 ? umma_desc(sbase + K::OFF_W1X, 128, (K::KP3 / 16) * 128, 0)
                                           : umma_desc(sbase + K::OFF_W1, 128, K::KSTEPS * 256, 0);
            const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32), b_lo0 = (uint32_t)db0;
            for (int n = 0; n < ntr; n++) {
                const int st = n % K::SX, ab = n & 1;
                TWAIT(0, sbase + K::BAR_FULL_X + 8 * st, (n / K::SX) & 1);
                TWAIT(1, sbase + K::BAR_T1_EMPTY + 8 * ab, ((n >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t a_lo = (uint32_t)da0 + st * (K::XS_BYTES >> 4);
                const uint32_t td = tmem_base + K::TM_L1 + ab * K::TM_L1_TRIAL;
                if (elect_one()) {
                    if (WOLF == 3) {
                        // super-rows: per tile of 128 super-rows KS3 k-steps of 32 bytes (two 16-byte chunks, 2048 B apart); N = 16 WS
#pragma unroll
                        for (int j = 0; j < K::TILES3; j++)
#pragma unroll
                            for (int ks = 0; ks < K::KS3; ks++)
                                umma_i8_lohi(td + 16 * K::WS * j, a_lo + ((j * K::TILE3 + ks * 4096) >> 4), a_hi, b_lo0 + ((ks * 256) >> 4),
                                             b_hi, idesc, ks > 0);
                    } else {
#pragma unroll
                        for (int j = 0; j < K::TILES; j++)
#pragma unroll
                            for (int ks = 0; ks < K::KSTEPS; ks++)
                                umma_i8_lohi(td + 16 * j, a_lo + ((j * K::TILE_BYTES + ks * (WOLF == 2 ? 32 : 4096)) >> 4), a_hi, b_lo0 + ((ks * 256) >> 4),
                                             b_hi, idesc, ks > 0);
                    }
                    umma_commit(sbase + K::BAR_EMPTY_X + 8 * st);
                    umma_commit(sbase + K::BAR_T1_FULL + 8 * ab);
                }
                __syncwarp();
            }
        }
    } else if (warp == kTcWarpM2) {
        // =====================================================================================
        // M2: layer-2 MMAs of a quad: per channel 3 x (M = 128 rows, N = 16 outputs, K = 32 samples)
        // =====================================================================================
        {
            constexpr uint32_t idesc = umma_idesc_i8_kk(128, 16);
            const uint64_t da0 = umma_desc(sbase + K::OFF_Y1, 16, 128, 0), db0 = umma_desc(sbase + K::OFF_W2, 128, 768, 0);
            const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32), b_lo0 = (uint32_t)db0;
            for (int gq = 0; gq < nq; gq++) {
                const int buf = gq & 1;
                TWAIT(0, sbase + K::BAR_Y1_FULL + 8 * buf, (gq >> 1) & 1);
                tc_fence_after();
                const uint32_t a_lo = (uint32_t)da0 + buf * ((4 * K::SLOT) >> 4);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    TWAIT(1, sbase + K::BAR_T2_EMPTY + 8 * h, (gq & 1) ^ 1);
                    tc_fence_after();
                    if (elect_one()) {
#pragma unroll
                        for (int k = 8 * h; k < 8 * h + 8; k++)
#pragma unroll
                            for (int ks = 0; ks < 3; ks++)
                                umma_i8_lohi(tmem_base + K::TM_L2 + 16 * k, a_lo + ((k * K::CHS + 32 * ks) >> 4), a_hi,
                                             b_lo0 + ((k * K::W2_CH + 256 * ks) >> 4), b_hi, idesc, ks > 0);
                        umma_commit(sbase + K::BAR_T2_FULL + 8 * h);
                        if (h == 1) umma_commit(sbase + K::BAR_Y1_EMPTY + 8 * buf);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp < kTcAW + kTcBW) {
        // =====================================================================================
        // B: layer-2 epilogue of (trial slot = warp & 3, channels 8 h .. 8 h + 7), then layer 3 of two channels
        // =====================================================================================
        const int wb = warp - kTcAW, s = warp & 3, h = wb >> 2;
        const int4* c2 = reinterpret_cast<const int4*>(smem + K::OFF_C2);
        for (int gq = 0; gq < nq; gq++) {
            const int buf = gq & 1, par = (gq >> 1) & 1;
            const long long b0 = 4 * (firstq + (long long)gq * strideq);
            // ---- layer 2: thread = row `lane` of slot s: conv outputs 16 lane .. 16 lane + 15 of 8 channels
            TWAIT(0, sbase + K::BAR_T2_FULL + 8 * h, gq & 1);
            tc_fence_after();
            TWAIT(1, sbase + K::BAR_Y3_EMPTY + 8 * buf, par ^ 1);      // the C warps have consumed y2[buf] of quad gq - 2
            uint8_t* y2s = smem + K::OFF_Y2 + buf * K::Y2_BUF + s * (16 * K::Y2S) + 8 + 2 * lane;
            // four channels per batch: 4 x 16 accumulators in flight, the TMEM half is released as soon as the
            // last batch sits in registers
#pragma unroll
            for (int kb = 0; kb < 2; kb++) {
                uint32_t r[4][16];
#ifdef MIBMI_PROF
                const long long tl0_ = clock64();
#endif
#pragma unroll
                for (int c = 0; c < 4; c++)
                    tmem_ld_32x32b_x16(r[c], tmem_base + ((uint32_t)(32 * s) << 16) + K::TM_L2 + 16 * (8 * h + 4 * kb + c));
                tmem_ld_wait();
#ifdef MIBMI_PROF
                pw_[2] += clock64() - tl0_;       // experiment: TMEM load latency, reported in the y3_empty column
#endif
                if (kb == 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(sbase + K::BAR_T2_EMPTY + 8 * h);
                }
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const int k = 8 * h + 4 * kb + c;
                    const int4 ca = c2[2 * k];                       // thr, sig, soff, m
                    const int shk = reinterpret_cast<const int*>(smem + K::OFF_C2)[8 * k + 4];
                    int s0 = 0, s1 = 0;
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        if (PLAIN) {
                            s0 += min(max(tdiv_magic((int)r[c][i] * ca.y + ca.z, ca.w, shk), 0), 127);
                            s1 += min(max(tdiv_magic((int)r[c][8 + i] * ca.y + ca.z, ca.w, shk), 0), 127);
                        } else {
                            s0 += max((int)r[c][i], ca.x);
                            s1 += max((int)r[c][8 + i], ca.x);
                        }
                    }
                    const uint32_t o01 = PLAIN ? (uint32_t)((s0 >> 3) | ((s1 >> 3) << 8))          // 0 .. 127 each
                                               : pack2_sat(tdiv_magic(s0 * ca.y + ca.z, ca.w, shk), tdiv_magic(s1 * ca.y + ca.z, ca.w, shk));
                    if (lane < K::ROWS) {
                        *reinterpret_cast<uint16_t*>(y2s + k * K::Y2S) = (uint16_t)o01;
                        if (TAPS && tap2 && b0 + s < B) {
                            int8_t* o = tap2 + ((size_t)(b0 + s) * 16 + k) * K::T8 + 2 * lane;
                            o[0] = (int8_t)(o01 & 0xff);
                            o[1] = (int8_t)((o01 >> 8) & 0xff);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + K::BAR_Y2_FULL + 8 * buf);
        }
    } else if (warp < kTcAW + kTcBW + kTcCW) {
        // =====================================================================================
        // C: layers 4 + 5 of the trial in slot (warp & 3) of every quad
        // =====================================================================================
        const int sl = warp & 3, cj = warp - (kTcAW + kTcBW);
        const uint32_t w4a0 = net->w4frag[0][lane], w4a1 = net->w4frag[1][lane];
        const int thr_lo = net->thr4[g], thr_hi = net->thr4[g + 8];
        const int sig_lo = net->sig4[g], sig_hi = net->sig4[g + 8];
        const int soff_lo = net->soff4[g], soff_hi = net->soff4[g + 8];
        const int m4_lo = net->m4[g], m4_hi = net->m4[g + 8], sh4_lo = net->sh4[g], sh4_hi = net->sh4[g + 8];
        const int b5 = net->b5[g & 3], m5 = net->m5, sh5 = net->sh5;
        const uint32_t l4_off = (lane & 15) * K::Y3S + 16 * (lane >> 4);
        uint8_t* y4 = smem + K::OFF_Y4 + cj * K::Y4_BYTES;
        uint32_t w5a[K::KST5][2];
#pragma unroll
        for (int ks = 0; ks < K::KST5; ks++) { w5a[ks][0] = net->w5frag[ks][0][lane]; w5a[ks][1] = net->w5frag[ks][1][lane]; }
        const uint32_t* y4w = reinterpret_cast<const uint32_t*>(y4);
        // layer 3: channels 4 cj .. 4 cj + 3 of the whole quad; MMA row r = 4 slot + block of 16 samples
        const int lrow = (lane & 7) + 8 * ((lane >> 3) & 1);
        uint32_t b3[4][2][2];
#pragma unroll
        for (int c = 0; c < 4; c++)
#pragma unroll
            for (int v = 0; v < 2; v++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++) b3[c][v][hh] = net->w3frag[4 * cj + c][v * 2 + hh][lane];
        const int m3 = net->m3, sh3 = net->sh3;
        // fused all-gather (NVLink P2P): the 4 x 4 logit bytes of a quad go to every peer's buffer as ONE 16-byte store
        // per peer (lane = peer), issued one quad late by the first C warp from the staging bytes of the four warps
        auto flush_peers = [&](int gprev) {
            const long long bq = 4 * (firstq + (long long)gprev * strideq);
            if (lane < peers.n) {
                int8_t* dst = peers.p[lane] + (size_t)bq * 4;
                const uint8_t* src = smem + K::OFF_LG + 16 * (gprev & 1);
                if (bq + 4 <= B && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
                    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
                } else {
                    for (int i = 0; i < 16 && bq * 4 + i < B * 4; i++) dst[i] = (int8_t)src[i];
                }
            }
        };
        for (int gq = 0; gq < nq; gq++) {
            const int buf = gq & 1, par = (gq >> 1) & 1;
            const long long b0 = 4 * (firstq + (long long)gq * strideq);
            const long long b = b0 + sl;
            TWAIT(0, sbase + K::BAR_Y2_FULL + 8 * buf, par);
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int k = 4 * cj + c;
                uint32_t a[4];
                ldsm_x4(a, sbase + K::OFF_Y2 + buf * K::Y2_BUF + ((lrow >> 2) * 16 + k) * K::Y2S + 16 * (lrow & 3) + 16 * (lane >> 4));
                int c3[2][4];
#pragma unroll
                for (int v = 0; v < 2; v++) {
                    c3[v][0] = c3[v][1] = c3[v][2] = c3[v][3] = 0;
                    mma_k32(c3[v], a, b3[c][v][0], b3[c][v][1]);
                }
#pragma unroll
                for (int hf = 0; hf < 2; hf++) {
                    const int row = g + 8 * hf, rs = row >> 2, i = row & 3;          // trial slot, block of 16 samples
                    uint8_t* y3b = smem + K::OFF_Y3 + (4 * buf + rs) * K::Y3_TRIAL + k * K::Y3S + 16 * (q >> 1) + 8 * (q & 1) +
                                   32 * (i >> 1) + 4 * (i & 1);
#pragma unroll
                    for (int v = 0; v < 2; v++) {
                        const int t = 16 * i + 8 * v + 2 * q;
                        const uint32_t w = FAST1 ? pack2_sat(tc_div1<true>(c3[v][2 * hf], m3, 0), tc_div1<true>(c3[v][2 * hf + 1], m3, 0))      // FAST1: layers 1 and 3 without the shift
                                                  : pack2_sat(tdiv_magic(c3[v][2 * hf], m3, sh3), tdiv_magic(c3[v][2 * hf + 1], m3, sh3));
                        if (t < K::T8) {
                            *reinterpret_cast<uint16_t*>(y3b + 2 * v) = (uint16_t)w;
                            if (TAPS && tap3 && b0 + rs < B) {
                                int8_t* o = tap3 + ((size_t)(b0 + rs) * 16 + k) * K::T8 + t;
                                o[0] = (int8_t)(w & 0xff);
                                o[1] = (int8_t)((w >> 8) & 0xff);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + K::BAR_Y3_EMPTY + 8 * buf);      // y2[buf] is free again
            asm volatile("bar.sync 1, %0;" ::"n"(kTcCW * 32) : "memory");        // y3[buf] of all 16 channels is complete
            if (peers.n > 0 && cj == 0 && gq > 0) flush_peers(gq - 1);            // ... and so are the staged logits of the previous quad
            const uint32_t y3a = sbase + K::OFF_Y3 + (4 * buf + sl) * K::Y3_TRIAL + l4_off;
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
                    if (PLAIN) {
                        s_lo += max(tdiv_magic(c[v][0] * sig_lo + soff_lo, m4_lo, sh4_lo), 0) + max(tdiv_magic(c[v][1] * sig_lo + soff_lo, m4_lo, sh4_lo), 0);
                        s_hi += max(tdiv_magic(c[v][2] * sig_hi + soff_hi, m4_hi, sh4_hi), 0) + max(tdiv_magic(c[v][3] * sig_hi + soff_hi, m4_hi, sh4_hi), 0);
                    } else {
                        s_lo += max(c[v][0], thr_lo) + max(c[v][1], thr_lo);
                        s_hi += max(c[v][2], thr_hi) + max(c[v][3], thr_hi);
                    }
                }
                const int grp = 4 * s4 + q;
                const int o_lo = PLAIN ? clip8(s_lo >> 3) : clip8(tdiv_magic(s_lo * sig_lo + soff_lo, m4_lo, sh4_lo));
                const int o_hi = PLAIN ? clip8(s_hi >> 3) : clip8(tdiv_magic(s_hi * sig_hi + soff_hi, m4_hi, sh4_hi));
                if (grp < K::T64) {
                    y4[g * K::T64P + grp] = (uint8_t)o_lo;
                    y4[(g + 8) * K::T64P + grp] = (uint8_t)o_hi;
                    if (TAPS && tap4 && b < B) {
                        tap4[((size_t)b * 16 + g) * K::T64 + grp] = (int8_t)o_lo;
                        tap4[((size_t)b * 16 + g + 8) * K::T64 + grp] = (int8_t)o_hi;
                    }
                }
            }
            __syncwarp();
            int z[4] = {0, 0, 0, 0};
#pragma unroll
            for (int ks = 0; ks < K::KST5; ks++) {
                const uint32_t a[4] = {w5a[ks][0], 0u, w5a[ks][1], 0u};
                const uint32_t bb0 = g == 0 ? y4w[8 * ks + q] : 0u, bb1 = g == 0 ? y4w[8 * ks + 4 + q] : 0u;
                mma_k32(z, a, bb0, bb1);
            }
            if (q == 0 && g < 4 && b < B) {
                const int8_t o = (int8_t)clip8(tdiv_magic(z[0] + b5, m5, sh5));
                logits[(size_t)b * 4 + g] = o;
                if (peers.n > 0) smem[K::OFF_LG + 16 * (gq & 1) + 4 * sl + g] = (uint8_t)o;
            }
            __syncwarp();
        }
        if (peers.n > 0 && nq > 0) {
            asm volatile("bar.sync 1, %0;" ::"n"(kTcCW * 32) : "memory");
            if (cj == 0) flush_peers(nq - 1);
        }
    }
    TPROF_END
    tc_fence_before();
    __syncthreads();
    if (warp == kTcWarpM1) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------------
template <int C, int T, int WOLF>
static cudaError_t launch_tc_instance(const PipeNet* net_dev, bool fast1, const int8_t* x, long long B, int8_t* logits,
                                      const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                                      cudaStream_t stream, bool plain = false) {
    using K = TCfg<C, T>;
    tmap_encode_t enc = tmap_encoder();
    if (!enc) return cudaErrorNotSupported;
    CUtensorMap tm;
    CUresult r;
    if (WOLF == 3) {
        // [B*T][C_ALIGN] seen as super-rows of WS samples = RB3 bytes (a multiple of 16): (16 bytes, super-rows, 16-byte chunks);
        // one box = [NCH3 chunks][128 super-rows][16 B]
        static_assert(WOLF != 3 || K::WS > 1, "super-row form is for rows that are not a multiple of 16 bytes");
        cuuint64_t dims[3] = {16, (cuuint64_t)(B * T / K::WS), (cuuint64_t)K::NCH3};
        cuuint64_t strides[2] = {(cuuint64_t)K::RB3, 16};
        cuuint32_t box[3] = {16, 128, (cuuint32_t)K::NCH3};
        cuuint32_t estr[3] = {1, 1, 1};
        r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<int8_t*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else if (WOLF == 2) {
        static_assert(WOLF != 2 || (C + 3) / 4 * 4 == 64, "64-byte rows");
        cuuint64_t dims[2] = {64, (cuuint64_t)(B * T)};
        cuuint64_t strides[1] = {64};
        cuuint32_t box[2] = {64, 128};
        cuuint32_t estr[2] = {1, 1};
        r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<int8_t*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else if (WOLF == 1) {
        // [B*T][C_ALIGN] seen as (16 bytes, time rows, 16-byte channel chunks): one box = [KPAD/16 chunks][128 rows][16 B]
        constexpr int CA = (C + 3) / 4 * 4;
        static_assert(WOLF != 1 || CA % 16 == 0, "Wolf-order rows must be a multiple of 16 bytes for this tensor map (else WOLF = 3)");
        cuuint64_t dims[3] = {16, (cuuint64_t)(B * T), (cuuint64_t)(CA / 16)};
        cuuint64_t strides[2] = {(cuuint64_t)CA, 16};
        cuuint32_t box[3] = {16, 128, (cuuint32_t)(K::KPAD / 16)};
        cuuint32_t estr[3] = {1, 1, 1};
        r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<int8_t*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        cuuint64_t dims[2] = {(cuuint64_t)T, (cuuint64_t)(B * C)};
        cuuint64_t strides[1] = {(cuuint64_t)T};
        cuuint32_t box[2] = {128, (cuuint32_t)C};
        cuuint32_t estr[2] = {1, 1};
        r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<int8_t*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    const bool taps = t1 || t2 || t3 || t4;
    if (plain && WOLF != 0) return cudaErrorNotSupported;      // the plain variant is instantiated for the module order only
    constexpr bool P0 = WOLF == 0;                             // (keeps the Wolf-order instantiations from doubling)
    auto kern = plain ? (taps ? tc_forward_kernel<K, false, true, WOLF, P0> : tc_forward_kernel<K, false, false, WOLF, P0>)
                : taps ? tc_forward_kernel<K, false, true, WOLF>
                       : (fast1 ? tc_forward_kernel<K, true, false, WOLF> : tc_forward_kernel<K, false, false, WOLF>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
    if (e != cudaSuccess) return e;
    long long grid = num_sms;
    const long long quads = (B + 3) / 4;
    if (grid > quads) grid = quads;
    kern<<<(unsigned)grid, kTcThreads, K::SMEM, stream>>>(tm, net_dev, B, logits, peers, t1, t2, t3, t4);
    return cudaGetLastError();
}

#define MIBMI_TC_SHAPES(X) X(64, 480) X(38, 480) X(24, 480) X(19, 480) X(16, 480) X(8, 480)

bool tc_has_instance(int C, int T) {
#define X(c, t) if (C == c && T == t) return true;
    MIBMI_TC_SHAPES(X)
#undef X
    return false;
}

cudaError_t launch_tc(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B, int8_t* logits,
                      const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream, bool plain) {
#define X(c, t) if (C == c && T == t) return launch_tc_instance<c, t, 0>(net_dev, fast1, x, B, logits, peers, t1, t2, t3, t4, num_sms, stream, plain);
    MIBMI_TC_SHAPES(X)
#undef X
    return cudaErrorInvalidValue;
}

#ifndef MIBMI_WOLF_NO_SW64
#define MIBMI_WOLF_NO_SW64 0
#endif
// Wolf order [B][T][C_ALIGN]: rows of 64 bytes (WOLF = 2), of 16 bytes (WOLF = 1), and -- WOLF = 3 -- rows that are not a multiple of
// 16 bytes, as super-rows of 2 or 4 samples (C_ALIGN = 40, 24, 20, 8)
#define MIBMI_TC_WOLF_SHAPES(X) X(64, 480) X(16, 480) X(38, 480) X(24, 480) X(19, 480) X(8, 480)

bool tc_has_wolf_instance(int C, int T) {
#define X(c, t) if (C == c && T == t) return true;
    MIBMI_TC_WOLF_SHAPES(X)
#undef X
    return false;
}

cudaError_t launch_tc_wolf(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B, int8_t* logits,
                           const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream) {
#define X(c, t) if (C == c && T == t) return launch_tc_instance<c, t, ((c + 3) / 4 * 4 == 64 && !MIBMI_WOLF_NO_SW64) ? 2 : (((c + 3) / 4 * 4) % 16 == 0 ? 1 : 3)>(net_dev, fast1, x, B, logits, peers, t1, t2, t3, t4, num_sms, stream);
    MIBMI_TC_WOLF_SHAPES(X)
#undef X
    return cudaErrorInvalidValue;
}

}  // namespace mibmi

#ifdef MIBMI_PROF
extern "C" __attribute__((visibility("default"))) int mibmi_debug_prof_tc(long long* out_host) {
    return (int)cudaMemcpyFromSymbol(out_host, mibmi::g_prof_tc, sizeof(mibmi::g_prof_tc));
}
#endif
