// Shared device-side definitions of the int8 edgeEEGNet / MI-BMInet path (sm_100a).
//
// Reference semantics (edge-eegnet_wolf/src/cl/net/layer1..5.c,
// python_utils/golden_model.py:153-379): int8 x int8 -> int32 accumulate, requantisation
// (acc + offset) / factor with C truncating division, clip to [-128, 127].
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mibmi {

constexpr int kNf = 64;       // temporal filter taps (golden_model.py:216-219)
constexpr int kL3Taps = 16;   // layer-3 taps
constexpr int kMaxF2 = 32;
constexpr int kMaxC = 256;
constexpr int kMaxN = 8;

// ---------------------------------------------------------------------------------------------
// Network as the generic (reference-order) kernel reads it.  Lives in global memory.
struct GenericNet {
    int C, T, F2, N, T8, T64, reorder;
    int fac1[kMaxF2], off1[kMaxF2];
    int fac2[kMaxF2], off2[kMaxF2];
    int fac4[kMaxF2], off4[kMaxF2];
    int fac3, fac5;
    int b5[kMaxN];
    int8_t w1[kMaxF2 * kMaxC];        // [F2][C]
    int8_t w2[kMaxF2 * kNf];          // [F2][64]  torch (cross-correlation) order
    int8_t w3[kMaxF2 * kL3Taps];      // [F2][16]  torch order (exported array un-flipped)
    int8_t w4[kMaxF2 * kMaxF2];       // [F2 out][F2 in]
    const int8_t* w5;                 // [N][F2][T64]  device pointer (size depends on T)
};

// ---------------------------------------------------------------------------------------------
// Network as the warp-specialised pipeline kernel (mibmi_pipe.cu) reads it: every array is already
// in MMA-fragment order, one 32-bit word per lane, so each warp loads its constants once with
// coalesced global loads and keeps them in registers for the whole launch (nothing of the network
// lives in shared memory).  Divisors are made positive on the host (the sign is folded into the L1 / L3 / L5 weights and
// into `sig` for the pooled layers) and replaced by (magic, shift) pairs:
//   tdiv(x, d) == (mulhi(x, magic) >> shift) + (x < 0)      for |x| < 2^29, d >= 4
constexpr int kPipeMaxKB16 = 4;     // layer-1 blocks of 16 channels (C <= 64)
constexpr int kPipeMaxKst5 = 16;    // layer-5 k-steps of 32 (T/64 padded to even <= 32)
constexpr int kW2U32 = 1792;        // bytes per channel of PipeNet::w2umma32
constexpr int kW1X = 64 * 96;       // PipeNet::w1x: N <= 64 (4 samples x 16 filters) x K <= 96 bytes

struct PipeNet {
    // layer 1: W1 as MMA B fragments per 16-channel block: [block][filter n-tile (0-7, 8-15)][lane]
    uint32_t w1blk[kPipeMaxKB16][2][32];
    int off1[16], m1[16], sh1[16];
    int fast1;                      // 1: sh1 == 0 everywhere (magic = floor(2^32/d)+1, exact after clip8)
    int plain;                      // 1: the plain (REORDER_BN off) variant, layer2.c:139-210 / layer4.c:99-140: the pooled layers
                                    // requantise EVERY accumulator, u = tdiv(v + (off >> 3), fac >> 3), then ReLU, sum of 8, >> 3;
                                    // sig / soff / m / sh of layers 2 and 4 then describe fac >> 3 and off >> 3 (thr unused)
    int pad0_[2];
    // layer 1 for tcgen05.mma: W1 [16 filters][KPAD channels] as K-major 8 x 16-byte core matrices,
    // byte (n >> 3) * (KPAD / 16) * 128 + (c >> 4) * 128 + (n & 7) * 16 + (c & 15), natural channel order
    alignas(16) uint8_t w1umma[16 * 64];
    // layer 2 for tcgen05.mma (mibmi_tc.cu): per channel the Toeplitz operand B[n][kk] = tap (kk - n - 1) of the
    // 64 taps, [16 outputs][96 samples] as K-major core matrices: byte (n >> 3) * 768 + (kk >> 4) * 128 + (n & 7) * 16 + (kk & 15)
    alignas(16) uint8_t w2umma[16 * 1536];
    // layer 2: Toeplitz B fragments [channel][variant*6 + kstep*2 + half][lane]
    uint32_t w2frag[16][24][32];
    int thr2[16], sig2[16], soff2[16], m2[16], sh2[16];
    // layer 3: Toeplitz B fragments [channel][variant*2 + half][lane]
    uint32_t w3frag[16][4][32];
    int m3, sh3, pad1_[2];
    // layer 4: A fragments [row half][lane]
    uint32_t w4frag[2][32];
    int thr4[16], sig4[16], soff4[16], m4[16], sh4[16];
    // layer 5: A fragments [kstep][k half][lane] (rows >= 4 are zero)
    uint32_t w5frag[kPipeMaxKst5][2][32];
    int b5[4];
    int m5, sh5, pad2_[2];
    // layer 2 for tcgen05.mma with N = 32 outputs per row (mibmi_tcr.cu): Toeplitz operand with REVERSED output order,
    // B'[n'][kk] = tap (kk + n' - 32) of the 64 taps (column n' = output 31 - n' of the row), n' < 32, kk < 96.  Reversal makes
    // B'[n' + 16][kk - 16] == B'[n'][kk], so the 8 x 16-byte core matrices may overlap: byte (n' >> 3) * 128 + (kk >> 4) * 256 +
    // (n' & 7) * 16 + (kk & 15) (descriptor LBO = 256 B, SBO = 128 B), 1792 bytes per channel instead of 3072
    alignas(16) uint8_t w2umma32[16 * kW2U32];
    // layer 1 for tcgen05.mma when the Wolf-order input [B][T][C_ALIGN] has rows that are not a multiple of 16 bytes (C_ALIGN = 40, 24,
    // 20, 8): S = 16 / gcd(C_ALIGN, 16) samples form one "super-row" of S * C_ALIGN bytes (a multiple of 16, so a tensor map can
    // describe it) and the weights become block structured: W1x[16 p + f][C_ALIGN p + c] = W1[f][c], p < S, zero elsewhere -- one MMA row
    // yields S samples x 16 filters (N = 16 S).  K-major core matrices over KP = ceil(S C_ALIGN / 32) * 32 bytes:
    // byte (n >> 3) * (KP / 16) * 128 + (k >> 4) * 128 + (n & 7) * 16 + (k & 15)
    alignas(16) uint8_t w1x[kW1X];
};
static_assert(sizeof(PipeNet) % 16 == 0, "PipeNet alignment");

// ---------------------------------------------------------------------------------------------
// Network as the sliding-window streaming kernel (mibmi_stream.cu) reads it.  Windows of T samples every
// 16 samples of a continuous stream share layer 1 and the interior of layer 2; what is window specific
// (the zero padding at the window edges) is folded into masked Toeplitz fragments:
//   w2L / w2R   layer-2 outputs 0..31 / T-32..T-1 of a window from its first / last 64 samples only
//   w3s         layer 3 with the window as the MMA row: class 0 / 1 / 2 = first / middle / last 16-output
//               block (inputs p < 4 and p > T/8 - 5 masked: they come from w3e instead)
//   w3e         layer-3 contribution of the 8 window-specific pooled samples (p = 0..3, T/8-4..T/8-1)
struct StreamNet {
    uint32_t w1nat[kPipeMaxKB16][2][32];   // W1 B fragments, natural channel order (channel = 16 block + row)
    int off1[16], m1[16], sh1[16];
    int fast1, pad0_[3];
    uint32_t w2frag[16][24][32];           // interior Toeplitz (same as PipeNet)
    uint32_t w2L[16][16][32];              // [channel][variant*4 + kstep*2 + half][lane]
    uint32_t w2R[16][16][32];
    int thr2[16], sig2[16], soff2[16], m2[16], sh2[16];
    uint32_t w3s[16][12][32];              // [channel][class*4 + e*2 + half][lane]
    uint32_t w3e[16][4][32];               // [channel][n-tile 0, 1, last-1, last][lane]  (k16 B fragments)
    int m3, sh3, pad1_[2];
    uint32_t w4frag[2][32];
    int thr4[16], sig4[16], soff4[16], m4[16], sh4[16];
    uint32_t w5frag[kPipeMaxKst5][2][32];
    int b5[4];
    int m5, sh5, pad2_[2];
};
static_assert(sizeof(StreamNet) % 16 == 0, "StreamNet alignment");

// C '/' : truncate toward zero.  (The generic kernel uses the hardware sequence as is.)
__device__ __forceinline__ int clip8(int x) { return min(max(x, -128), 127); }

__device__ __forceinline__ int tdiv_magic(int x, int m, int s) {
    return (__mulhi(x, m) >> s) + (int)((unsigned)x >> 31);
}

}  // namespace mibmi
