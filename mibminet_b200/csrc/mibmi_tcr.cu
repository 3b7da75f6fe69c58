// tcgen05 / TMEM forward for trials whose rows are NOT 16-byte aligned and whose length is not a multiple of 32
// (22 x 1125, BASELINE config 3).  Same tensor-core formulation as mibmi_tc.cu (layer 1 = tile x W1, layer 2 = Hankel view of
// the y1 byte array x Toeplitz taps, accumulators in TMEM), with these differences:
//   * a trial starts at an arbitrary byte, so TMA tiles are impossible: the producer bulk-copies the trial from the
//     16-byte boundary below it into a raw stage and four REALIGN warps build the layer-1 operand tiles:
//       WOLF = 0  x [B][C][T] (module order): every row is funnel-shifted into the 128-byte-swizzled MN-major tile layout the TMA
//                 unit would have produced;
//       WOLF = 1  x [B][T][C_ALIGN] (the order net_model_compute receives, gen_input_header.py:75-76): a sample is 24 contiguous
//                 bytes (8-byte aligned) = already K-major for layer 1; the warps only expand it to two 16-byte chunks (channels
//                 0-15, 16-31) of a no-swizzle K-major tile -- no shifts, no scratch, no second pass over HBM;
//   * layer 2 computes N = 32 outputs per MMA row: the y1 buffer is read with a row pitch of 32 bytes through a K-major
//     SWIZZLE_32B descriptor (the bytes are stored where the absolute-address XOR of that mode puts them; k-step ks = start + 32 ks,
//     i.e. "the next row": the Hankel view) against a Toeplitz operand with reversed output order whose core matrices overlap
//     (PipeNet::w2umma32; both verified on the box with tools/probe_umma3.cu).  A trial takes 37 rows, a group of G = 3 trials
//     = 111 rows = ONE tile of 128 rows: 16 channels x 3 k-steps = 48 instructions per group (96 with N = 16);
//   * layers 3-5 of a trial run in ONE warp (private y3 / y4), as T/8 = 140 needs 9 blocks of 16 samples.
// One persistent CTA per SM, 31 warps:
//   A  0-11  layer-1 epilogue (3 warps per TMEM lane quarter)      layer1.c:96-104
//   B 12-19  layer-2 epilogue (2 warps per lane quarter)           layer2.c:99-118
//   C 20-23  layers 3 + 4 + 5 of one trial each                    layer3.c:70, layer4.c:99-140, layer5.c:81-89
//   R 24-27  raw bytes -> layer-1 operand tiles                    (rt_dma_memcpy L2 -> L1, layer1.c:137-154)
//   P 28     bulk-copy producer     M1 29  layer-1 MMAs (layer1.c:87-96)     M2 30  layer-2 MMAs (layer2.c:87-118)
// Bit-identical to the reference (tests/test_parity_gpu.py).
#include <cuda.h>
#include <string.h>

#include "mibmi_common.cuh"
#include "mibmi_kernels.h"
#include "mibmi_ptx.cuh"

namespace mibmi {

#ifndef MIBMI_TCR_SLEEP_NS
#define MIBMI_TCR_SLEEP_NS 64
#endif
#ifndef MIBMI_TCR_SLEEP_LONG_NS
#define MIBMI_TCR_SLEEP_LONG_NS 128       // waits with a trial's worth of slack (producer, M1, free-buffer waits)
#endif
#ifndef MIBMI_TCR_WAIT
#define MIBMI_TCR_WAIT 1                  // mbar_wait_mode: 0 = test_wait + nanosleep, 1 = test_wait then try_wait (r2a: +3 %), 2 = try_wait
#endif
#ifdef MIBMI_PROF
__device__ long long g_prof_tcr[148 * 32 * 4];
#define RWAIT(slot, ...) do { if (!mbar_test(__VA_ARGS__)) { const long long t0_ = clock64(); mbar_wait_mode<MIBMI_TCR_WAIT>(__VA_ARGS__, MIBMI_TCR_SLEEP_NS); pw_[slot] += clock64() - t0_; } } while (0)
#define RWAITL(slot, ...) do { if (!mbar_test(__VA_ARGS__)) { const long long t0_ = clock64(); mbar_wait_mode<MIBMI_TCR_WAIT>(__VA_ARGS__, MIBMI_TCR_SLEEP_LONG_NS); pw_[slot] += clock64() - t0_; } } while (0)
#define RPROF_BEGIN long long pw_[3] = {0, 0, 0}; const long long pt0_ = clock64();
#define RPROF_END if ((threadIdx.x & 31) == 0) { long long* o_ = g_prof_tcr + (blockIdx.x * 32 + (threadIdx.x >> 5)) * 4; \
    o_[0] = clock64() - pt0_; o_[1] = pw_[0]; o_[2] = pw_[1]; o_[3] = pw_[2]; }
#else
#define RWAIT(slot, ...) mbar_wait_mode<MIBMI_TCR_WAIT>(__VA_ARGS__, MIBMI_TCR_SLEEP_NS)
#define RWAITL(slot, ...) mbar_wait_mode<MIBMI_TCR_WAIT>(__VA_ARGS__, MIBMI_TCR_SLEEP_LONG_NS)
#define RPROF_BEGIN
#define RPROF_END
#endif

#ifndef MIBMI_TCR_L3_UNROLL
#define MIBMI_TCR_L3_UNROLL 4
#endif
#ifndef MIBMI_TCR_L4_UNROLL
#define MIBMI_TCR_L4_UNROLL 1
#endif
constexpr int kTcrL3Unroll = MIBMI_TCR_L3_UNROLL, kTcrL4Unroll = MIBMI_TCR_L4_UNROLL;   // layers 3 / 4 loops of the C warps
#ifndef MIBMI_TCR_RW
#define MIBMI_TCR_RW 4
#endif
#ifndef MIBMI_TCR_AW
#define MIBMI_TCR_AW 12                   // layer-1 epilogue warps (a multiple of 4: one per TMEM lane quarter)
#endif
#ifndef MIBMI_TCR_HELP
#define MIBMI_TCR_HELP 1                  // the first set of layer-1 epilogue warps is a third consumer of layer-2 batches (third TMEM half)
#endif                                    // while it would otherwise wait for tile_done: the phases A / M2+B alternate on the single y1 buffer
constexpr int kRAW = MIBMI_TCR_AW, kRNA = kRAW / 4, kRBW = 8, kRCW = 4, kRRW = MIBMI_TCR_RW;
constexpr int kRL2Sets = 2 + (MIBMI_TCR_HELP ? 1 : 0);                        // consumer sets (4 warps each) = TMEM halves of layer 2
__host__ __device__ constexpr int tcr_l2_batches(int set) { return (8 - set + kRL2Sets - 1) / kRL2Sets; }   // batches bt = set, set + kRL2Sets, ... < 8
static_assert(kRAW % 4 == 0 && kRAW >= 4, "layer-1 epilogue warps come in sets of four");
constexpr int kRWarpB = kRAW, kRWarpC = kRAW + kRBW, kRWarpR = kRWarpC + kRCW, kRWarpP = kRWarpR + kRRW,
              kRWarpM1 = kRWarpP + 1, kRWarpM2 = kRWarpP + 2;
#ifndef MIBMI_TCR_ROW2
#define MIBMI_TCR_ROW2 0                  // 1: realign warps keep the loads of two rows in flight (needs SETREG; r2o2: slower, 75.7 vs 84 M trials/s)
#endif
#ifndef MIBMI_TCR_SETREG
#define MIBMI_TCR_SETREG 0                // 1 (r2o3: 89.8 vs 88.6 M trials/s in Wolf order, 83.1 vs 84.5 M in module order: off) = 32 warps: the producer / MMA warpgroup (+ one idle warp) shrinks to kRRegsPM registers, the realign warpgroup grows to
#endif                                    // kRRegsR (setmaxnreg) and keeps the loads of two rows in flight
constexpr bool kRSetReg = MIBMI_TCR_SETREG && (kRWarpM2 + 2 == 32) && kRWarpR % 4 == 0 && kRRW == 4;
constexpr int kRRegsPM = 40, kRRegsR = 88;            // 64 - 24 and 64 + 24: four warps each
constexpr int kRThreads = (kRWarpM2 + 1 + (MIBMI_TCR_SETREG ? 1 : 0)) * 32;
static_assert(!MIBMI_TCR_SETREG || kRSetReg, "setmaxnreg needs whole warpgroups: 12 A + 8 B + 4 C warps, then R = warps 24-27, then P, M1, M2, idle");

__host__ __device__ constexpr int tcr_pitch16(int n) { return n % 4 == 2 ? n : tcr_pitch16(n + 1); }   // smallest p >= n, p = 2 (mod 4)

template <int C_, int T_, int WOLF_>
struct RCfg {
    static constexpr int C = C_, T = T_, WOLF = WOLF_, G = 3;
    static constexpr int CA = (C + 3) / 4 * 4;                                  // bytes per sample in Wolf order
    static constexpr int TRIAL_BYTES = WOLF ? T * CA : C * T;
    static constexpr int TILES = (T + 127) / 128, UNITS = (T + 31) / 32, CHUNKS = (T + 15) / 16;
    // layer-1 operand tiles.  WOLF = 0: MN-major [C][128 samples], 128-byte swizzle, rows 24..31 of a tile may alias the next
    // tile (zero weights); the whole trial per stage.  WOLF = 1: K-major [2 chunks][128 samples][16 B]; half a trial per stage.
    static constexpr int TILE_PITCH = WOLF ? 4096 : (C <= 24 ? 3072 : 4096);
    static constexpr int PARTS = WOLF ? 2 : 1;                                  // hand-offs R -> M1 per trial
    static constexpr int PART0_TILES = WOLF ? 4 : TILES;                        // tiles of part 0 (part 1: the rest)
    static constexpr int PART_TILES_MAX = WOLF ? (TILES - 4 > 4 ? TILES - 4 : 4) : TILES;
    static constexpr int XS_BYTES = ((PART_TILES_MAX * TILE_PITCH + 1023) / 1024) * 1024;
    static constexpr int RAW_BYTES = ((TRIAL_BYTES + 15 + 48 + 127) / 128) * 128;   // + misalignment + the realign warps' read-ahead
    static constexpr int T8 = T / 8, T64 = T8 / 8;
    static constexpr int VROWS = (8 * T8) / 32;                               // layer-2 rows (32 conv outputs) per trial
    static constexpr int SLOT = ((32 + T + 31) / 32) * 32, SROWS = SLOT / 32;   // y1 bytes / rows per trial: 32 zeros | T | zeros
    static constexpr int GROWS = G * SROWS;
    // y1 row of one filter.  stmatrix writes 8 rows x 16 B per phase at chunk (a >> 4) ^ ((a >> 7) & 1) (the 32-byte swizzle of the
    // layer-2 operand): the eight rows fall into eight different 16-byte bank groups iff pitch / 16 = 2 (mod 4).  (An odd pitch / 16,
    // the rule for an unswizzled layout, cost 6 wavefronts per stmatrix.x4 instead of 4: ncu r2e, 72 excess wavefronts per trial.)
    static constexpr int CHS = 16 * tcr_pitch16((G * SLOT + 15) / 16);
    static constexpr int NW3 = (T8 + 15) / 16, Y2S = 16 * (NW3 + 1), Y2_TRIAL = 16 * Y2S;
    static constexpr int Y3CH = 2 * ((NW3 + 1) / 2), Y3S = 16 * (Y3CH + 1), Y3_BYTES = 16 * Y3S;
    static constexpr int NST = (T64 + 3) / 4, T64P = (T64 + 1) / 2 * 2, KST5 = T64P / 2;
    static constexpr int Y4_BYTES = ((16 * T64P + 15) / 16) * 16;
    // TMEM columns: layer-1 accumulators of two trials, then kRL2Sets layer-2 "halves" of 2 channels x 32 outputs
    static constexpr int TM_L1_TRIAL = 16 * TILES, TM_L2 = 2 * TM_L1_TRIAL, TM_L2_HALF = 64;
    // barriers
    static constexpr int BAR_RAW_FULL = 0, BAR_RAW_EMPTY = 16, BAR_X_FULL = 32, BAR_X_EMPTY = 48, BAR_T1_FULL = 64,
                         BAR_T1_EMPTY = 80, BAR_Y1_FULL = 96 /* G */, BAR_TILE_DONE = 128, BAR_T2_FULL = 136 /* kRL2Sets */,
                         BAR_T2_EMPTY = 160 /* kRL2Sets */, BAR_Y2_FULL = 192 /* 2 */, BAR_Y2_EMPTY = 232, BAR_END = 248;
    // shared memory map
    static constexpr int OFF_TMEM = 248, OFF_C2 = 256, OFF_W1 = 1024, OFF_W2 = 2048;
    static constexpr int OFF_XS = ((OFF_W2 + 16 * kW2U32 + 1023) / 1024) * 1024;
    static constexpr int OFF_RAW = OFF_XS + 2 * XS_BYTES;
    static constexpr int OFF_Y1 = OFF_RAW + 2 * RAW_BYTES;
    static constexpr int OFF_Y2 = OFF_Y1 + 16 * CHS;
    static constexpr int OFF_Y3 = OFF_Y2 + 2 * G * Y2_TRIAL;
    static constexpr int OFF_Y4 = OFF_Y3 + kRCW * Y3_BYTES;
    static constexpr int OFF_W3 = ((OFF_Y4 + kRCW * Y4_BYTES + 127) / 128) * 128;      // w3frag [16][4][32] words
    static constexpr int OFF_W5 = OFF_W3 + 16 * 4 * 32 * 4;                             // w5frag [KST5][2][32] words
    static constexpr int SMEM = ((OFF_W5 + KST5 * 2 * 32 * 4 + 127) / 128) * 128;
    static_assert(C <= 32 && CA <= 32, "one k-step of 32 channels");
    static_assert(!WOLF || CA == 24, "Wolf-order expansion is written for rows of 24 bytes");
    static_assert((8 * T8) % 32 == 0 && T8 % 4 == 0, "whole layer-2 rows of 32 outputs");
    static_assert(SLOT - 32 - T >= 8 * T8 + 32 - T, "right padding of layer 2 inside the slot");
    static_assert(GROWS <= 128, "a group is one MMA tile");
    static_assert((CHS / 16) % 4 == 2 && CHS >= G * SLOT, "conflict-free stmatrix under the 32-byte swizzle");
    static_assert(G <= 4, "barrier slots");
    static_assert(TM_L2 + kRL2Sets * TM_L2_HALF <= 512 && kRL2Sets <= 3, "TMEM map / barrier slots");
    static_assert(NW3 <= 16 && KST5 <= kPipeMaxKst5, "layer 3 / 5 limits");
    static_assert(XS_BYTES % 1024 == 0 && OFF_XS % 1024 == 0, "swizzle atoms are 1024-byte aligned");
    static_assert(SMEM <= 227 * 1024, "shared memory");
    __host__ __device__ static constexpr int part_first(int h) { return h == 0 ? 0 : PART0_TILES; }
    __host__ __device__ static constexpr int part_tiles(int h) { return PARTS == 1 ? TILES : (h == 0 ? PART0_TILES : TILES - PART0_TILES); }
};

template <bool FAST>
__device__ __forceinline__ int tcr_div1(int x, int m, int s) {
    if (FAST) return __mulhi(x, m) + (int)((unsigned)x >> 31);
    return (__mulhi(x, m) >> s) + (int)((unsigned)x >> 31);
}
// layer 1 with the BN offset folded into the multiplication (FAST form only): hi32((x + off) m) = hi32(x m + off m) -- one
// multiply-add with a 64-bit constant instead of an add and a multiply-high; the quotient's sign bit stands in for (x + off < 0)
// (m > 0, so hi32 is negative exactly when x + off is).  |x + off| < 2^29 and m < 2^31 keep the 64-bit sum exact.
#ifndef MIBMI_TCR_FOLD
#define MIBMI_TCR_FOLD 1
#endif
__device__ __forceinline__ int tcr_div1_fold(int x, int m, long long offm) {
    const int hi = (int)(((long long)x * m + offm) >> 32);
    return hi + (int)((unsigned)hi >> 31);
}
__device__ __forceinline__ void tcr_tmem_ld_32x32b_x16(uint32_t (&r)[16], uint32_t taddr) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr));
}
__device__ __forceinline__ constexpr uint32_t tcr_idesc_i8_kk(int M, int N) {   // both operands K-major
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint4 tcr_lds128(const uint8_t* p) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
}
__device__ __forceinline__ uint2 tcr_lds64(const uint8_t* p) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(smem_u32(p)));
    return v;
}
// WOLF = 0: one row of T samples: lane = 16-byte chunk (+ 32 per pass); all loads of the row are issued before the shifts
#ifndef MIBMI_TCR_TAIL
#define MIBMI_TCR_TAIL 0                  // 1: the <= 8 chunks beyond the second pass of every row are realigned four rows at a time
                                          // (13 instead of 16.5 passes per warp and trial; r2j: bit-exact but slower, 82.8 vs 83.7 M trials/s -- the
                                          // per-lane word selects and the bank conflicts of four rows in one pass cost what the passes save)
#endif
template <class K>
struct TcrTail {
    static constexpr int TAILC = K::CHUNKS - 64;                                 // chunks of a row beyond two full passes (7 at T = 1125)
    static constexpr bool SPLIT = MIBMI_TCR_TAIL && TAILC > 0 && TAILC <= 8;
    static constexpr int RPT = SPLIT ? 32 / TAILC : 1;                           // rows per tail pass
};
template <class K>
struct TcrRow {                                                                   // the aligned 32 bytes around every chunk of a row (lane = chunk)
    static constexpr int P = TcrTail<K>::SPLIT ? 2 : (K::CHUNKS + 31) / 32;
    uint4 q0[P], q1[P];
};
template <class K>
__device__ __forceinline__ void tcr_row_load(TcrRow<K>& r, const uint8_t* rbase, int lane) {
#pragma unroll
    for (int p = 0; p < TcrRow<K>::P; p++)
        if (32 * p + 31 < K::CHUNKS || lane + 32 * p < K::CHUNKS) {
            r.q0[p] = tcr_lds128(rbase + 512 * p);          // kept 128 bits wide: the narrower loads the compiler would pick for
            r.q1[p] = tcr_lds128(rbase + 512 * p + 16);     // the 5 words in use are 4-way bank conflicted at this lane stride
        }
}
template <class K, int K4>
__device__ __forceinline__ void tcr_row_store(const TcrRow<K>& r, uint8_t* xrow, int c, int sh, int lane) {
#pragma unroll
    for (int p = 0; p < TcrRow<K>::P; p++) {
        const int jj = lane + 32 * p;
        if (32 * p + 31 < K::CHUNKS || jj < K::CHUNKS) {
            const uint32_t w[8] = {r.q0[p].x, r.q0[p].y, r.q0[p].z, r.q0[p].w, r.q1[p].x, r.q1[p].y, r.q1[p].z, r.q1[p].w};
            uint4 o;
            o.x = __funnelshift_r(w[K4], w[K4 + 1], sh);
            o.y = __funnelshift_r(w[K4 + 1], w[K4 + 2], sh);
            o.z = __funnelshift_r(w[K4 + 2], w[K4 + 3], sh);
            o.w = __funnelshift_r(w[K4 + 3], w[K4 + 4], sh);
            *reinterpret_cast<uint4*>(xrow + (jj >> 3) * K::TILE_PITCH + (((jj & 7) ^ (c & 7)) << 4)) = o;
        }
    }
}
// WOLF = 0: one row of T samples: lane = 16-byte chunk (+ 32 per pass); all loads of the row are issued before the shifts
template <class K>
__device__ __forceinline__ void tcr_row_shift(const TcrRow<K>& r, uint8_t* xs, int c, int src0, int lane) {
    const int sh = 8 * (src0 & 3);                         // warp-uniform byte shift, word shift below
    uint8_t* xrow = xs + c * 128;
    switch ((src0 >> 2) & 3) {
        case 0: tcr_row_store<K, 0>(r, xrow, c, sh, lane); break;
        case 1: tcr_row_store<K, 1>(r, xrow, c, sh, lane); break;
        case 2: tcr_row_store<K, 2>(r, xrow, c, sh, lane); break;
        default: tcr_row_store<K, 3>(r, xrow, c, sh, lane); break;
    }
}

// PLAIN: the REORDER_BN-off variant of the pooled layers (see mibmi_tc.cu)
template <class K, bool FAST1, bool TAPS, bool PLAIN = false>
__global__ void __launch_bounds__(kRThreads, 1)
tcr_forward_kernel(const int8_t* __restrict__ x, const PipeNet* __restrict__ net, long long B, int8_t* __restrict__ logits,
                   const PeerOut peers, int8_t* __restrict__ tap1, int8_t* __restrict__ tap2, int8_t* __restrict__ tap3,
                   int8_t* __restrict__ tap4) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    const uint32_t sbase = smem_u32(smem);
    const long long ngroups_total = (B + K::G - 1) / K::G;
    const long long firstg = blockIdx.x, strideg = gridDim.x;
    const int ng = firstg < ngroups_total ? (int)((ngroups_total - firstg + strideg - 1) / strideg) : 0;   // groups of this CTA
    const int ntr = K::G * ng;                                                                              // trial slots
    auto trial_of = [&](int n) -> long long { return K::G * (firstg + (long long)(n / K::G) * strideg) + (n % K::G); };

    // ---- prologue ------------------------------------------------------------------------------------
    if (tid == 0) {
        for (int i = 0; i < 2; i++) {
            mbar_init(sbase + K::BAR_RAW_FULL + 8 * i, 1);
            mbar_init(sbase + K::BAR_RAW_EMPTY + 8 * i, kRRW);
            mbar_init(sbase + K::BAR_X_FULL + 8 * i, kRRW);
            mbar_init(sbase + K::BAR_X_EMPTY + 8 * i, 1);            // commit of M1
            mbar_init(sbase + K::BAR_T1_FULL + 8 * i, 1);            // commit of M1
            mbar_init(sbase + K::BAR_T1_EMPTY + 8 * i, kRAW);
            mbar_init(sbase + K::BAR_Y2_EMPTY + 8 * i, K::G);
            mbar_init(sbase + K::BAR_Y2_FULL + 8 * i, 4 * kRL2Sets);
        }
        for (int i = 0; i < kRL2Sets; i++) {
            mbar_init(sbase + K::BAR_T2_FULL + 8 * i, 1);            // commit of M2
            mbar_init(sbase + K::BAR_T2_EMPTY + 8 * i, 4);           // the four warps (TMEM lane quarters) of the consumer set
        }
        for (int j = 0; j < K::G; j++) mbar_init(sbase + K::BAR_Y1_FULL + 8 * j, kRAW);
        mbar_init(sbase + K::BAR_TILE_DONE, 1);                      // commit of M2: the y1 buffer of the group has been read
        mbar_fence_init();
    }
    for (int i = tid; i < 16 * 32 / 4; i += kRThreads)
        reinterpret_cast<uint32_t*>(smem + K::OFF_W1)[i] = reinterpret_cast<const uint32_t*>(net->w1umma)[i];
    for (int i = tid; i < 16 * kW2U32 / 16; i += kRThreads)
        reinterpret_cast<uint4*>(smem + K::OFF_W2)[i] = reinterpret_cast<const uint4*>(net->w2umma32)[i];
    for (int i = tid; i < 16 * 8; i += kRThreads) {
        const int k = i >> 3, f = i & 7;
        reinterpret_cast<int*>(smem + K::OFF_C2)[i] = f == 0 ? net->thr2[k] : f == 1 ? net->sig2[k] : f == 2 ? net->soff2[k]
                                                      : f == 3 ? net->m2[k] : f == 4 ? net->sh2[k] : 0;
    }
    {
        uint4* z = reinterpret_cast<uint4*>(smem + K::OFF_XS);     // tiles (rows >= C are read with zero weights), y1 gaps, y2 / y3 padding
        for (int i = tid; i < (K::OFF_W3 - K::OFF_XS) / 16; i += kRThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    for (int i = tid; i < 16 * 4 * 32; i += kRThreads)
        reinterpret_cast<uint32_t*>(smem + K::OFF_W3)[i] = (&net->w3frag[0][0][0])[i];
    for (int i = tid; i < K::KST5 * 2 * 32; i += kRThreads)
        reinterpret_cast<uint32_t*>(smem + K::OFF_W5)[i] = (&net->w5frag[0][0][0])[i];
    if (warp == kRWarpM1) tmem_alloc(sbase + K::OFF_TMEM, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<const uint32_t*>(smem + K::OFF_TMEM);
    RPROF_BEGIN

    // Layer-2 epilogue of one consumer set (four warps, one per TMEM lane quarter): the batches bt = set, set + kRL2Sets, ... of
    // group gq from TMEM half `set`.  Thread = TMEM lane = row 32 (warp & 3) + lane of the group's y1 buffer = 32 consecutive conv
    // outputs (four pooling groups) of trial row / SROWS; channels 2 bt, 2 bt + 1.  Column n' of the accumulator = output 31 - n' of
    // the row (reversed Toeplitz, PipeNet::w2umma32).  ws0 / ws1: wait-profile slots of the calling role
    auto l2_consume = [&](const int set, const int gq, const int ws0, const int ws1) {
        (void)ws0; (void)ws1;
        const int qi = warp & 3;
        const int4* c2 = reinterpret_cast<const int4*>(smem + K::OFF_C2);
        const int row = 32 * qi + lane;
        const int j = row / K::SROWS, blk = row - j * K::SROWS;
        const bool valid = j < K::G && blk < K::VROWS;
        const int gpar = gq & 1;
        const int cnt = tcr_l2_batches(set);
        RWAITL(ws1, sbase + K::BAR_Y2_EMPTY + 8 * gpar, ((gq >> 1) & 1) ^ 1);     // the C warps have consumed y2[gpar] of group gq - 2
        const long long b = K::G * (firstg + (long long)gq * strideg) + j;
        uint8_t* y2s = smem + K::OFF_Y2 + (gpar * K::G + (j < K::G ? j : 0)) * K::Y2_TRIAL + 8 + 4 * blk;
#pragma unroll 1
        for (int i = 0; i < cnt; i++) {
            const int bt = set + kRL2Sets * i;
            const int use = gq * cnt + i;                                   // how often this half has been used before
            RWAIT(ws0, sbase + K::BAR_T2_FULL + 8 * set, use & 1);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < 2; c++) {
                uint32_t r[2][16];
                const uint32_t ta = tmem_base + ((uint32_t)(32 * qi) << 16) + K::TM_L2 + K::TM_L2_HALF * set + 32 * c;
                tcr_tmem_ld_32x32b_x16(r[0], ta);
                tcr_tmem_ld_32x32b_x16(r[1], ta + 16);
                tmem_ld_wait();
                if (c == 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(sbase + K::BAR_T2_EMPTY + 8 * set);
                }
                const int k = 2 * bt + c;
                const int4 ca = c2[2 * k];                       // thr, sig, soff, m
                const int shk = reinterpret_cast<const int*>(smem + K::OFF_C2)[8 * k + 4];
                int s[4];                                        // s[p] = pooling group p = outputs 8 p .. 8 p + 7 = columns 24 - 8 p .. 31 - 8 p
#pragma unroll
                for (int p = 0; p < 4; p++) {
                    const int col = 24 - 8 * p;
                    int acc = 0;
#pragma unroll
                    for (int e = 0; e < 8; e++)
                        acc += PLAIN ? min(max(tdiv_magic((int)r[col >> 4][(col & 15) + e] * ca.y + ca.z, ca.w, shk), 0), 127)
                                     : max((int)r[col >> 4][(col & 15) + e], ca.x);
                    s[p] = PLAIN ? (acc >> 3) : tdiv_magic(acc * ca.y + ca.z, ca.w, shk);
                }
                const uint32_t o = pack4_sat(s[0], s[1], s[2], s[3]);
                if (valid) {
                    *reinterpret_cast<uint32_t*>(y2s + k * K::Y2S) = o;
                    if (TAPS && tap2 && b < B) *reinterpret_cast<uint32_t*>(tap2 + ((size_t)b * 16 + k) * K::T8 + 4 * blk) = o;
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(sbase + K::BAR_Y2_FULL + 8 * gpar);
    };

    if (warp < kRAW) {
        // =====================================================================================
        // A: layer-1 epilogue.  TMEM lanes 32 (warp & 3) .. + 31 of the tiles jt = (warp >> 2) + n (mod kRNA), + kRNA, ...
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
        const int qi = warp & 3, a3 = warp >> 2;
        for (int n = 0; n < ntr; n++) {
            const int gq = n / K::G, j = n - gq * K::G, ab = n & 1;
            RWAIT(0, sbase + K::BAR_T1_FULL + 8 * ab, (n >> 1) & 1);
            tc_fence_after();
            const uint32_t yb = st_off + K::SLOT * j;
            const long long b = trial_of(n);
            // the warp's tiles jt0, jt0 + 3, ...: the TMEM load of the next unit is in flight while this one is requantised
            const int jt0 = (a3 + n) % kRNA;
            const uint32_t ta0 = tmem_base + ((uint32_t)(32 * qi) << 16) + ab * K::TM_L1_TRIAL;
            uint32_t r[2][2][8];
            if (32 * (4 * jt0 + qi) < K::T) {
                tmem_ld_16x256b_x2(r[0][0], ta0 + 16 * jt0);
                tmem_ld_16x256b_x2(r[0][1], ta0 + 16 * jt0 + (16u << 16));
            }
            // the y1 buffer is single: the layer-2 MMAs of the previous group must have read it (one commit per group)
            RWAITL(1, sbase + K::BAR_TILE_DONE, (gq & 1) ^ 1);
#pragma unroll
            for (int it = 0; it < (K::TILES + kRNA - 1) / kRNA; it++) {
                const int jt = jt0 + kRNA * it, u = 4 * jt + qi;
                if (jt >= K::TILES || 32 * u >= K::T) break;
                tmem_ld_wait();
                if (it + 1 < (K::TILES + kRNA - 1) / kRNA && jt + kRNA < K::TILES && 32 * (u + 4 * kRNA) < K::T) {
                    tmem_ld_16x256b_x2(r[(it + 1) & 1][0], ta0 + 16 * (jt + kRNA));
                    tmem_ld_16x256b_x2(r[(it + 1) & 1][1], ta0 + 16 * (jt + kRNA) + (16u << 16));
                }
                uint32_t w[2][2];
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int nt = 0; nt < 2; nt++) {
                        int v[4];
#pragma unroll
                        for (int c = 0; c < 4; c++)
                            v[c] = (FAST1 && MIBMI_TCR_FOLD)
                                       ? tcr_div1_fold((int)r[it & 1][mt][4 * nt + c], mg1[nt][c & 1], offm[nt][c & 1])
                                       : tcr_div1<FAST1>((int)r[it & 1][mt][4 * nt + c] + off1[nt][c & 1], mg1[nt][c & 1], sh1[nt][c & 1]);
                        w[mt][nt] = pack4_sat(v[0], v[1], v[2], v[3]);
                        if (TAPS && tap1 && b < B)
#pragma unroll
                            for (int c = 0; c < 4; c++) {
                                const int f = 8 * nt + 2 * q + (c & 1), t = 32 * u + 16 * mt + g + 8 * (c >> 1);
                                if (t < K::T) tap1[((size_t)b * 16 + f) * K::T + t] = (int8_t)clip8(v[c]);
                            }
                    }
                if (32 * u + 32 > K::T) {                   // samples >= T are layer 2's zero padding (warp-uniform branch)
#pragma unroll
                    for (int mt = 0; mt < 2; mt++) {
                        const int lim = K::T - 32 * u - 16 * mt;       // bytes 0, 1 = sample g, bytes 2, 3 = sample g + 8
                        const uint32_t m = (g < lim ? 0x0000ffffu : 0u) | (g + 8 < lim ? 0xffff0000u : 0u);
                        w[mt][0] &= m;
                        w[mt][1] &= m;
                    }
                }
                // 32-byte swizzle of the layer-2 operand: 16-byte chunk bit ^= address bit 7 (absolute shared address)
                const uint32_t ya = yb + 32 * u;
                stsm_t8_x4(ya ^ ((ya >> 3) & 16u), w[0][0], w[0][1], w[1][0], w[1][1]);
            }
            tc_fence_before();
            fence_proxy_async();          // y1 is read by the tensor core (async proxy)
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(sbase + K::BAR_T1_EMPTY + 8 * ab);
                mbar_arrive(sbase + K::BAR_Y1_FULL + 8 * j);
            }
            // the group's y1 is complete: until tile_done the layer-1 epilogue has nothing to do -- the first set takes the layer-2
            // batches of the third TMEM half meanwhile
            if (MIBMI_TCR_HELP && a3 == 0 && j == K::G - 1) l2_consume(kRL2Sets - 1, gq, 2, 2);
        }
    } else if (warp >= kRWarpR && warp < kRWarpP) {
        // the realign warps are one warpgroup: they take the registers the producer / MMA warpgroup gives up (two rows in flight)
        if (MIBMI_TCR_SETREG) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRRegsR));
        // =====================================================================================
        // R: raw bytes -> layer-1 operand tiles
        // =====================================================================================
        const int rw = warp - kRWarpR;
        for (int n = 0; n < ntr; n++) {
            const int st = n & 1;
            RWAIT(0, sbase + K::BAR_RAW_FULL + 8 * st, (n >> 1) & 1);
            const long long b = trial_of(n);
            const int o15 = (int)(((unsigned long long)b * K::TRIAL_BYTES) & 15);
            const uint8_t* raw = smem + K::OFF_RAW + st * K::RAW_BYTES;
            if (K::WOLF) {
                // sample t = 24 bytes at raw + o15 + 24 t (8-byte aligned): chunk 0 = bytes 0..15, chunk 1 = bytes 16..31 (the
                // last 8 belong to the next sample and meet zero weights).  lane = sample: conflict-free 8-byte loads
                // (lane stride 24 B) and 16-byte stores (lane stride 16 B)
#pragma unroll
                for (int h = 0; h < K::PARTS; h++) {
                    RWAITL(1, sbase + K::BAR_X_EMPTY + 8 * h, (n & 1) ^ 1);
                    if (b < B) {
                        uint8_t* xs = smem + K::OFF_XS + h * K::XS_BYTES;
                        const int t0 = 128 * K::part_first(h), nit = 4 * K::part_tiles(h);      // warp iterations of 32 samples
                        for (int it = rw; it < nit; it += kRRW) {
                            const int tl = 32 * it + lane, t = t0 + tl;
                            if (t < K::T) {
                                const uint8_t* s = raw + o15 + 24 * t;
                                const uint2 w0 = tcr_lds64(s), w1 = tcr_lds64(s + 8), w2 = tcr_lds64(s + 16), w3 = tcr_lds64(s + 24);
                                uint8_t* d = xs + (tl >> 7) * K::TILE_PITCH + 16 * (tl & 127);
                                *reinterpret_cast<uint4*>(d) = make_uint4(w0.x, w0.y, w1.x, w1.y);
                                *reinterpret_cast<uint4*>(d + 2048) = make_uint4(w2.x, w2.y, w3.x, w3.y);
                            }
                        }
                    }
                    fence_proxy_async();          // the tiles are read by the tensor core (async proxy)
                    __syncwarp();
                    if (lane == 0) mbar_arrive(sbase + K::BAR_X_FULL + 8 * h);
                }
            } else {
                RWAITL(1, sbase + K::BAR_X_EMPTY + 8 * st, ((n >> 1) & 1) ^ 1);
                if (b < B) {
                    // [C][128 samples] tiles with the 128-byte swizzle (16-byte chunk j of row c at j ^ (c & 7))
                    uint8_t* xs = smem + K::OFF_XS + st * K::XS_BYTES;
                    if (kRSetReg && MIBMI_TCR_ROW2) {
                        // two rows in flight (88 registers): the loads of the next row are issued before the shifts of this one
                        TcrRow<K> ra, rb;
                        auto src_of = [&](int c) { return o15 + c * K::T; };
                        auto base_of = [&](int c) { return raw + (src_of(c) & ~15) + 16 * lane; };
                        tcr_row_load<K>(ra, base_of(rw), lane);
#pragma unroll 1
                        for (int c = rw; c < K::C; c += 2 * kRRW) {
                            const bool two = c + kRRW < K::C;
                            if (two) tcr_row_load<K>(rb, base_of(c + kRRW), lane);
                            tcr_row_shift<K>(ra, xs, c, src_of(c), lane);
                            if (two) {
                                if (c + 2 * kRRW < K::C) tcr_row_load<K>(ra, base_of(c + 2 * kRRW), lane);
                                tcr_row_shift<K>(rb, xs, c + kRRW, src_of(c + kRRW), lane);
                            }
                        }
                    } else {
                        for (int c = rw; c < K::C; c += kRRW) {
                            const int src0 = o15 + c * K::T;                       // first byte of the row in the raw stage
                            TcrRow<K> r;
                            tcr_row_load<K>(r, raw + (src0 & ~15) + 16 * lane, lane);
                            tcr_row_shift<K>(r, xs, c, src0, lane);
                        }
                    }
                    if (TcrTail<K>::SPLIT) {
                        // the chunks 64 .. of the warp's rows, RPT rows per pass: lane = (row, chunk), so shift and word offset are per
                        // lane: the five words are picked with two rounds of selects (word offset bits 1 and 0) before the funnel shifts
                        constexpr int TAILC = TcrTail<K>::TAILC, RPT = TcrTail<K>::RPT;
                        constexpr int NROWS = (K::C + kRRW - 1) / kRRW, NTP = (NROWS + RPT - 1) / RPT;
                        const int ri = lane / TAILC, jt = lane - ri * TAILC;
#pragma unroll
                        for (int tp = 0; tp < NTP; tp++) {
                            const int c = rw + kRRW * (RPT * tp + ri);
                            if (ri < RPT && c < K::C) {
                                const int src0 = o15 + c * K::T;
                                const uint8_t* a = raw + (src0 & ~15) + 16 * (64 + jt);
                                const uint4 q0 = tcr_lds128(a), q1 = tcr_lds128(a + 16);
                                const int k4 = (src0 >> 2) & 3, sh = 8 * (src0 & 3);
                                const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
                                uint32_t v[6], u[5];
#pragma unroll
                                for (int i = 0; i < 6; i++) v[i] = (k4 & 2) ? w[i + 2] : w[i];
#pragma unroll
                                for (int i = 0; i < 5; i++) u[i] = (k4 & 1) ? v[i + 1] : v[i];
                                uint4 o;
                                o.x = __funnelshift_r(u[0], u[1], sh);
                                o.y = __funnelshift_r(u[1], u[2], sh);
                                o.z = __funnelshift_r(u[2], u[3], sh);
                                o.w = __funnelshift_r(u[3], u[4], sh);
                                const int jj = 64 + jt;
                                *reinterpret_cast<uint4*>(xs + c * 128 + (jj >> 3) * K::TILE_PITCH + (((jj & 7) ^ (c & 7)) << 4)) = o;
                            }
                        }
                    }
                }
                fence_proxy_async();          // the tiles are read by the tensor core (async proxy)
                __syncwarp();
                if (lane == 0) mbar_arrive(sbase + K::BAR_X_FULL + 8 * st);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + K::BAR_RAW_EMPTY + 8 * st);
        }
    }
    else if (warp >= kRWarpP) {
        // P, M1, M2 (and the idle warp that completes the warpgroup) need few registers
        if (MIBMI_TCR_SETREG) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRRegsPM));
    if (warp == kRWarpP) {
        // =====================================================================================
        // P: one bulk copy per trial, from the 16-byte boundary below its first byte.  No byte outside the batch is
        //    read: the copy of the LAST trial of the batch stops at the 16-byte boundary below its end and the lanes
        //    fetch the remaining <= 15 bytes with byte loads
        // =====================================================================================
        const unsigned long long total = (unsigned long long)B * K::TRIAL_BYTES;
        for (int n = 0; n < ntr; n++) {
            const int st = n & 1;
            RWAITL(0, sbase + K::BAR_RAW_EMPTY + 8 * st, ((n >> 1) & 1) ^ 1);
            const long long b = trial_of(n);
            const uint32_t bar = sbase + K::BAR_RAW_FULL + 8 * st;
            if (b < B) {
                const unsigned long long o = (unsigned long long)b * K::TRIAL_BYTES;
                const unsigned long long a0 = o & ~15ull;
                unsigned long long a1 = (o + K::TRIAL_BYTES + 15) & ~15ull;
                if (a1 > total) {                                  // warp-uniform: only the last trial of the batch
                    a1 = total & ~15ull;
                    if (a1 + lane < total) smem[K::OFF_RAW + st * K::RAW_BYTES + (a1 - a0) + lane] = (uint8_t)x[a1 + lane];
                    __syncwarp();                                  // the byte stores are ordered before lane 0's arrive (release)
                }
                if (lane == 0) {
                    const uint32_t len = (uint32_t)(a1 - a0);
                    mbar_expect_tx(bar, len);
                    bulk_g2s(sbase + K::OFF_RAW + st * K::RAW_BYTES, x + a0, len, bar);
                }
            } else if (lane == 0) {
                mbar_arrive(bar);     // padding trial of the last group: nothing to read
            }
            __syncwarp();
        }
    } else if (warp == kRWarpM1) {
        // =====================================================================================
        // M1: layer-1 MMAs, one per tile (K = 32 channels covers C)
        // =====================================================================================
        constexpr uint32_t idesc = K::WOLF ? tcr_idesc_i8_kk(128, 16) : umma_idesc_i8(128, 16);
        const uint64_t da0 = K::WOLF ? umma_desc(sbase + K::OFF_XS, 2048, 128, 0) : umma_desc(sbase + K::OFF_XS, 16, 1024, 2);
        const uint64_t db0 = umma_desc(sbase + K::OFF_W1, 128, 256, 0);
        const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32), b_lo0 = (uint32_t)db0;
        for (int n = 0; n < ntr; n++) {
            const int ab = n & 1;
            const uint32_t td = tmem_base + ab * K::TM_L1_TRIAL;
#pragma unroll
            for (int h = 0; h < K::PARTS; h++) {
                const int st = K::PARTS == 1 ? (n & 1) : h;
                const int par = K::PARTS == 1 ? ((n >> 1) & 1) : (n & 1);
                RWAITL(0, sbase + K::BAR_X_FULL + 8 * st, par);
                if (h == 0) RWAITL(1, sbase + K::BAR_T1_EMPTY + 8 * ab, ((n >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t a_lo = (uint32_t)da0 + st * (K::XS_BYTES >> 4);
                if (elect_one()) {
#pragma unroll
                    for (int j = 0; j < K::part_tiles(h); j++)
                        umma_i8_lohi(td + 16 * (K::part_first(h) + j), a_lo + ((j * K::TILE_PITCH) >> 4), a_hi, b_lo0, b_hi, idesc, 0);
                    umma_commit(sbase + K::BAR_X_EMPTY + 8 * st);
                    if (h == K::PARTS - 1) umma_commit(sbase + K::BAR_T1_FULL + 8 * ab);
                }
                __syncwarp();
            }
        }
    } else if (warp == kRWarpM2) {
        // =====================================================================================
        // M2: layer-2 MMAs: one tile of 128 rows x 32 outputs per group; 8 batches of 2 channels x 3 k-steps into
        //     alternating TMEM halves.  A = y1 row of the channel, K-major SWIZZLE_32B (rows 32 B apart, SBO = 256 B), k-step ks
        //     starts 32 ks bytes further; B = w2umma32 of the channel (LBO = 256 B, SBO = 128 B), k-step ks = + 512 ks bytes
        // =====================================================================================
        constexpr uint32_t idesc = tcr_idesc_i8_kk(128, 32);
        const uint64_t da0 = umma_desc(sbase + K::OFF_Y1, 16, 256, 6), db0 = umma_desc(sbase + K::OFF_W2, 256, 128, 0);
        const uint32_t a_hi = (uint32_t)(da0 >> 32), b_hi = (uint32_t)(db0 >> 32), b_lo0_ = (uint32_t)db0, a_lo0_ = (uint32_t)da0;
        for (int gq = 0; gq < ng; gq++) {
#pragma unroll
            for (int j = 0; j < K::G; j++) RWAIT(0, sbase + K::BAR_Y1_FULL + 8 * j, gq & 1);
            tc_fence_after();
            // the descriptor words of the 48 instructions are immediates away from these two: keeping them opaque per group stops the
            // compiler from hoisting 96 precomputed words out of the group loop (they would not fit the 40 registers of this warpgroup)
            uint32_t a_lo0 = a_lo0_, b_lo0 = b_lo0_;
            asm volatile("" : "+r"(a_lo0), "+r"(b_lo0));
#pragma unroll
            for (int bt = 0; bt < 8; bt++) {
                const int h = bt % kRL2Sets;                                // TMEM half = consumer set
                const int use = gq * tcr_l2_batches(h) + bt / kRL2Sets;     // how often this half has been used before
                RWAIT(1, sbase + K::BAR_T2_EMPTY + 8 * h, (use & 1) ^ 1);
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int c = 0; c < 2; c++)
#pragma unroll
                        for (int ks = 0; ks < 3; ks++) {
                            const int k = 2 * bt + c;
                            umma_i8_lohi(tmem_base + K::TM_L2 + K::TM_L2_HALF * h + 32 * c,
                                         a_lo0 + ((k * K::CHS + 32 * ks) >> 4), a_hi,
                                         b_lo0 + ((k * kW2U32 + 512 * ks) >> 4), b_hi, idesc, ks > 0);
                        }
                    umma_commit(sbase + K::BAR_T2_FULL + 8 * h);
                    if (bt == 7) umma_commit(sbase + K::BAR_TILE_DONE);
                }
                __syncwarp();
            }
        }
    }
    } else if (warp >= kRWarpB && warp < kRWarpC) {
        // =====================================================================================
        // B: layer-2 epilogue, consumer sets 0 and 1 (l2_consume above)
        // =====================================================================================
        const int set = (warp - kRWarpB) >> 2;
        for (int gq = 0; gq < ng; gq++) l2_consume(set, gq, 0, 1);
    } else if (warp >= kRWarpC && warp < kRWarpR) {
        // =====================================================================================
        // C: layers 3 + 4 + 5 of the trials n = cj (mod 4); y3 / y4 are private to the warp
        // =====================================================================================
        const int cj = warp - kRWarpC;
        const uint32_t w4a0 = net->w4frag[0][lane], w4a1 = net->w4frag[1][lane];
        const int thr_lo = net->thr4[g], thr_hi = net->thr4[g + 8];
        const int sig_lo = net->sig4[g], sig_hi = net->sig4[g + 8];
        const int soff_lo = net->soff4[g], soff_hi = net->soff4[g + 8];
        const int m4_lo = net->m4[g], m4_hi = net->m4[g + 8], sh4_lo = net->sh4[g], sh4_hi = net->sh4[g + 8];
        const int b5 = net->b5[g & 3], m5 = net->m5, sh5 = net->sh5;
        const int m3 = net->m3, sh3 = net->sh3;
        // ldmatrix.x4 lane addressing: matrix j = lane >> 3, row (lane & 7) + 8 (j & 1), k-half j >> 1
        const int lrow = (lane & 7) + 8 * ((lane >> 3) & 1);
        const int l3_row = lrow < K::NW3 ? lrow : K::NW3 - 1;
        const uint32_t l3_off = 16 * (l3_row + (lane >> 4));
        uint8_t* y3 = smem + K::OFF_Y3 + cj * K::Y3_BYTES;
        uint8_t* y4 = smem + K::OFF_Y4 + cj * K::Y4_BYTES;
        const uint32_t* y4w = reinterpret_cast<const uint32_t*>(y4);
        const uint32_t* w3s = reinterpret_cast<const uint32_t*>(smem + K::OFF_W3) + lane;
        const uint32_t* w5s = reinterpret_cast<const uint32_t*>(smem + K::OFF_W5) + lane;
        // layer-3 stores: sample t = 16 i + 8 v + 2 q (+1) goes to chunk 2 (i >> 1) + (q >> 1), position
        // 8 (q & 1) + 4 (i & 1) + 2 v: the order layer 4 wants (pooling group = MMA column pair)
        const uint32_t y3_thr = 16 * (q >> 1) + 8 * (q & 1);
        const uint32_t y3a = smem_u32(y3) + (lane & 15) * K::Y3S + 16 * (lane >> 4);
        for (int n = cj; n < ntr; n += kRCW) {
            const int gq = n / K::G, j = n - gq * K::G, gpar = gq & 1;
            const long long b = trial_of(n);
            RWAIT(0, sbase + K::BAR_Y2_FULL + 8 * gpar, (gq >> 1) & 1);
            const uint32_t y2a = sbase + K::OFF_Y2 + (gpar * K::G + j) * K::Y2_TRIAL + l3_off;
#pragma unroll kTcrL3Unroll
            for (int k = 0; k < 16; k++) {
                uint32_t a[4];
                ldsm_x4(a, y2a + k * K::Y2S);
                int c3[2][4];
#pragma unroll
                for (int v = 0; v < 2; v++) {
                    c3[v][0] = c3[v][1] = c3[v][2] = c3[v][3] = 0;
                    mma_k32(c3[v], a, w3s[(4 * k + 2 * v) * 32], w3s[(4 * k + 2 * v + 1) * 32]);
                }
                uint8_t* y3b = y3 + k * K::Y3S + y3_thr;
#pragma unroll
                for (int half = 0; half < (K::NW3 > 8 ? 2 : 1); half++) {
                    const int i = g + 8 * half;
#pragma unroll
                    for (int v = 0; v < 2; v++) {
                        const int t = 16 * i + 8 * v + 2 * q;
                        const uint32_t w = FAST1 ? pack2_sat(tcr_div1<true>(c3[v][2 * half], m3, 0), tcr_div1<true>(c3[v][2 * half + 1], m3, 0))  // FAST1: layers 1 and 3 without the shift
                                                  : pack2_sat(tdiv_magic(c3[v][2 * half], m3, sh3), tdiv_magic(c3[v][2 * half + 1], m3, sh3));
                        if (t < K::T8) {
                            *reinterpret_cast<uint16_t*>(y3b + 32 * (i >> 1) + 4 * (i & 1) + 2 * v) = (uint16_t)w;
                            if (TAPS && tap3 && b < B) {
                                tap3[((size_t)b * 16 + k) * K::T8 + t] = (int8_t)(w & 0xff);
                                tap3[((size_t)b * 16 + k) * K::T8 + t + 1] = (int8_t)((w >> 8) & 0xff);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sbase + K::BAR_Y2_EMPTY + 8 * gpar);      // this trial's y2 buffer is free again
#pragma unroll kTcrL4Unroll
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
            // ---- layer 5: D[class][0] = W5[class][k, t] . y4[k, t]; only MMA column 0 carries data
            int z[4] = {0, 0, 0, 0};
#pragma unroll
            for (int ks = 0; ks < K::KST5; ks++) {
                const uint32_t a[4] = {w5s[(2 * ks) * 32], 0u, w5s[(2 * ks + 1) * 32], 0u};
                const uint32_t bb0 = g == 0 ? y4w[8 * ks + q] : 0u, bb1 = g == 0 ? y4w[8 * ks + 4 + q] : 0u;
                mma_k32(z, a, bb0, bb1);
            }
            // the four logit bytes of the trial (lanes g = 0..3, q = 0) become one word: one 4-byte store to the local buffer and
            // one per peer (fused gather over NVLink P2P; lane 0 = local, lane i = peer i - 1)
            const uint32_t ob = (q == 0 && g < 4) ? ((uint32_t)(uint8_t)(int8_t)clip8(tdiv_magic(z[0] + b5, m5, sh5)) << (8 * g)) : 0u;
            const uint32_t word = __reduce_or_sync(0xffffffffu, ob);
            if (b < B && lane <= peers.n) {
                int8_t* dst = lane == 0 ? logits : peers.p[lane - 1];
                *reinterpret_cast<uint32_t*>(dst + (size_t)b * 4) = word;
            }
            __syncwarp();   // y3 / y4 are rewritten by the warp's next trial
        }
    }
    RPROF_END
    tc_fence_before();
    __syncthreads();
    if (warp == kRWarpM1) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------------
template <int C, int T, int WOLF>
static cudaError_t launch_tcr_instance(const PipeNet* net_dev, bool fast1, const int8_t* x, long long B, int8_t* logits,
                                       const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                                       cudaStream_t stream, bool plain) {
    using K = RCfg<C, T, WOLF>;
    const bool taps = t1 || t2 || t3 || t4;
    if (plain && WOLF != 0) return cudaErrorNotSupported;      // the plain variant is instantiated for the module order only
    constexpr bool P0 = WOLF == 0;
    auto kern = plain ? (taps ? tcr_forward_kernel<K, false, true, P0> : tcr_forward_kernel<K, false, false, P0>)
                : taps ? tcr_forward_kernel<K, false, true>
                       : (fast1 ? tcr_forward_kernel<K, true, false> : tcr_forward_kernel<K, false, false>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
    if (e != cudaSuccess) return e;
    long long grid = num_sms;
    const long long groups = (B + K::G - 1) / K::G;
    if (grid > groups) grid = groups;
    kern<<<(unsigned)grid, kRThreads, K::SMEM, stream>>>(x, net_dev, B, logits, peers, t1, t2, t3, t4);
    return cudaGetLastError();
}

#define MIBMI_TCR_SHAPES(X) X(22, 1125)

bool tcr_has_instance(int C, int T) {
#define X(c, t) if (C == c && T == t) return true;
    MIBMI_TCR_SHAPES(X)
#undef X
    return false;
}
bool tcr_has_wolf_instance(int C, int T) { return tcr_has_instance(C, T); }

cudaError_t launch_tcr(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B, int layout, int8_t* logits,
                       const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream, bool plain) {
#define X(c, t)                                                                                                                     \
    if (C == c && T == t)                                                                                                           \
        return layout == 0 ? launch_tcr_instance<c, t, 0>(net_dev, fast1, x, B, logits, peers, t1, t2, t3, t4, num_sms, stream, plain)     \
                           : launch_tcr_instance<c, t, 1>(net_dev, fast1, x, B, logits, peers, t1, t2, t3, t4, num_sms, stream, plain);
    MIBMI_TCR_SHAPES(X)
#undef X
    return cudaErrorInvalidValue;
}

}  // namespace mibmi

#ifdef MIBMI_PROF
extern "C" __attribute__((visibility("default"))) int mibmi_debug_prof_tcr(long long* out_host) {
    return (int)cudaMemcpyFromSymbol(out_host, mibmi::g_prof_tcr, sizeof(mibmi::g_prof_tcr));
}
#endif
