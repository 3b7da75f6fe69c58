// Fused int8 edgeEEGNet / MI-BMInet forward for sm_100a: one persistent CTA streams trials through
// all five layers; inter-layer activations never leave shared memory.
//
//   input    cp.async.bulk (TMA bulk copy, mbarrier complete_tx) of one [C][T] trial per stage,
//            double buffered
//   layer 1  spatial conv  D[16 filters x 8 times] += W1[16 x 32 ch] * X[32 ch x 8 times]
//            (mma.sync m16n8k32 s8; the [C][T] -> channel-contiguous byte transpose is done in
//            registers with PRMT), offset preloaded in the accumulator, magic-number division,
//            saturating pack, 64-bit stores of 8 consecutive samples per filter row
//            -- edge-eegnet_wolf/src/cl/net/layer1.c:87-96
//   layer 2  depthwise 64-tap temporal conv as a Hankel x Toeplitz product per channel:
//            A[16 windows x 32] (ldmatrix straight from the layer-1 rows, windows 32 samples
//            apart) times 4 Toeplitz variants B_v[32 x 8] of the taps; 3 k-steps cover the
//            32+63 samples a window needs.  Each thread ends up with complete pooling groups
//            of 8, so ReLU-threshold + sum-pool + requant need no shuffles
//            -- layer2.c:87-118 (REORDER_BN)
//   layer 3  depthwise 16-tap conv, dp4a against byte-phase-shifted taps -- layer3.c:70
//   layer 4  pointwise conv, mma.sync m16n8k16 s8, pool via 2 shuffles -- layer4.c:99-140
//   layer 5  classifier, dp4a + warp reduction -- layer5.c:81-89
//
// Integer results are bit-identical to the reference (tests/test_parity_gpu.py).
#include "mibmi_common.cuh"
#include "mibmi_kernels.h"
#include "mibmi_ptx.cuh"

namespace mibmi {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

template <int C_, int T_>
struct Cfg {
    static constexpr int C = C_, T = T_;
    static constexpr int KSTEPS = (C + 31) / 32;
    static constexpr int T8 = T / 8, T64 = T8 / 8;
    static constexpr int UNITS = (T + 31) / 32;          // layer-1 work units of 32 samples
    static constexpr int MROWS = (T8 * 8 + 31) / 32;     // layer-2 windows (32 conv outputs each)
    static constexpr int MT = (MROWS + 15) / 16;         // layer-2 m-tiles
    static constexpr int Y1A = 32 + 32 * UNITS, Y1B = 32 * (MROWS - 1) + 96;
    static constexpr int Y1NEED = Y1A > Y1B ? Y1A : Y1B;
    static constexpr int Y1S = ((Y1NEED - 32 + 127) / 128) * 128 + 32;   // = 32 mod 128: conflict-free stores
    static constexpr int Y2S = ((T8 + 24 + 15) / 16) * 16;
    static constexpr int T8R = (T8 + 7) / 8 * 8;
    static constexpr int T64P = (T64 + 3) / 4 * 4;
    static constexpr int XROWS = 32 * KSTEPS;
    static constexpr int XSTRIDE = T;                    // bytes; T % 16 == 0 on this path
    static constexpr int TRIAL_BYTES = C * T;
    static constexpr int XS_BYTES = ((XROWS * XSTRIDE + 128 + 127) / 128) * 128;
    // shared memory map
    static constexpr int OFF_BAR = 0;
    static constexpr int OFF_XS = 128;
    static constexpr int OFF_Y1 = OFF_XS + 2 * XS_BYTES;
    static constexpr int OFF_Y2 = OFF_Y1 + 16 * Y1S + 128;
    static constexpr int OFF_Y3 = OFF_Y2 + 16 * Y2S;
    static constexpr int OFF_Y4 = OFF_Y3 + T8R * 16;
    static constexpr int OFF_NET = ((OFF_Y4 + 16 * T64P + 15) / 16) * 16;
    static constexpr int SMEM = OFF_NET + (int)sizeof(FusedNet);
    static_assert(T % 16 == 0, "bulk-copy path needs 16-byte rows");
    static_assert(KSTEPS >= 1 && KSTEPS <= 2, "C <= 64");
    static_assert(T64P <= kMaxT64P, "T too large for FusedNet::w5p");
};

// ---------------------------------------------------------------------------------------------
template <class K>
__global__ void __launch_bounds__(kThreads, 2)
fused_forward_kernel(const FusedNet* __restrict__ net_g, const int8_t* __restrict__ x, long long B,
                     int8_t* __restrict__ logits) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;

    uint8_t* xs = smem + K::OFF_XS;
    uint8_t* y1 = smem + K::OFF_Y1;
    uint8_t* y2 = smem + K::OFF_Y2;
    uint8_t* y3 = smem + K::OFF_Y3;   // [T8R][16] time-major for the layer-4 B operand
    uint8_t* y4 = smem + K::OFF_Y4;   // [16][T64P]
    const FusedNet& net = *reinterpret_cast<const FusedNet*>(smem + K::OFF_NET);
    const uint32_t bar0 = smem_u32(smem + K::OFF_BAR);

    const long long first = blockIdx.x, stride = gridDim.x;

    // ---- prologue: barriers, first two copies, network -> smem, zero the padded activation rows
    if (tid == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar0 + 8, 1);
        mbar_fence_init();
        for (int s = 0; s < 2; s++) {
            const long long b = first + s * stride;
            if (b < B) {
                mbar_expect_tx(bar0 + 8 * s, K::TRIAL_BYTES);
                bulk_g2s(smem_u32(xs + s * K::XS_BYTES), x + (size_t)b * K::TRIAL_BYTES, K::TRIAL_BYTES, bar0 + 8 * s);
            }
        }
    }
    {
        const uint4* src = reinterpret_cast<const uint4*>(net_g);
        uint4* dst = reinterpret_cast<uint4*>(smem + K::OFF_NET);
        for (int i = tid; i < (int)(sizeof(FusedNet) / 16); i += kThreads) dst[i] = src[i];
        uint4* z = reinterpret_cast<uint4*>(smem + K::OFF_Y1);
        for (int i = tid; i < (K::OFF_NET - K::OFF_Y1) / 16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
        if (K::TRIAL_BYTES < K::XROWS * K::XSTRIDE + 128) {   // rows >= C are multiplied by zero weights;
            for (int s = 0; s < 2; s++) {                     // keep them defined anyway
                uint32_t* t = reinterpret_cast<uint32_t*>(xs + s * K::XS_BYTES + K::TRIAL_BYTES);
                for (int i = tid; i < (K::XS_BYTES - K::TRIAL_BYTES) / 4; i += kThreads) t[i] = 0;
            }
        }
    }
    __syncthreads();

    // ---- per-thread constants
    uint32_t afrag[K::KSTEPS][4];
#pragma unroll
    for (int s = 0; s < K::KSTEPS; s++) {
        afrag[s][0] = net.w1frag[((s * 2 + 0) * 16 + g) * 4 + q];
        afrag[s][1] = net.w1frag[((s * 2 + 0) * 16 + g + 8) * 4 + q];
        afrag[s][2] = net.w1frag[((s * 2 + 1) * 16 + g) * 4 + q];
        afrag[s][3] = net.w1frag[((s * 2 + 1) * 16 + g + 8) * 4 + q];
    }
    const int off1_lo = net.off1[g], off1_hi = net.off1[g + 8];
    const int m1_lo = net.m1[g], m1_hi = net.m1[g + 8], sh1_lo = net.sh1[g], sh1_hi = net.sh1[g + 8];
    // layer-2 Toeplitz fragment addressing (see host packer): variant v, this lane
    int toep_base[4];
#pragma unroll
    for (int v = 0; v < 4; v++) {
        const int o = 8 * (g >> 1) + 2 * v + (g & 1);
        const int j0 = 31 - o;                       // byte position of tap index -32 relative to kappa = 4q
        toep_base[v] = (j0 & 3) * kToepCopyStride + (j0 & ~3) + 4 * q;
    }

    long long b_prev = -1;
    int it = 0;
    for (long long b = first; b < B; b += stride, ++it) {
        const int st = it & 1;
        const uint32_t* xw = reinterpret_cast<const uint32_t*>(xs + st * K::XS_BYTES);

        // ---- layer 4 of the previous trial (its y3 is complete since the last barrier)
        if (b_prev >= 0) {
            for (int j = warp; j < K::T64; j += kWarps) {
                int acc[4] = {0, 0, 0, 0};
                const uint32_t bfrag = reinterpret_cast<const uint32_t*>(y3)[(8 * j + g) * 4 + q];
                mma_k16(acc, net.w4frag[g * 4 + q], net.w4frag[(g + 8) * 4 + q], bfrag);
                const int tl = net.thr4[g], th = net.thr4[g + 8];
                int s_lo = max(acc[0], tl) + max(acc[1], tl);
                int s_hi = max(acc[2], th) + max(acc[3], th);
                s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 1);
                s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 1);
                s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 2);
                s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 2);
                if (q == 0) {
                    const int xl = s_lo * net.sig4[g] + net.soff4[g];
                    const int xh = s_hi * net.sig4[g + 8] + net.soff4[g + 8];
                    y4[g * K::T64P + j] = (uint8_t)clip8(tdiv_magic(xl, net.m4[g], net.sh4[g]));
                    y4[(g + 8) * K::T64P + j] = (uint8_t)clip8(tdiv_magic(xh, net.m4[g + 8], net.sh4[g + 8]));
                }
            }
        }

        // ---- layer 1
        mbar_wait(bar0 + 8 * st, (it >> 1) & 1);
        for (int u = warp; u < K::UNITS; u += kWarps) {
            int acc[4][4];
#pragma unroll
            for (int j = 0; j < 4; j++) { acc[j][0] = acc[j][1] = off1_lo; acc[j][2] = acc[j][3] = off1_hi; }
#pragma unroll
            for (int s = 0; s < K::KSTEPS; s++) {
                uint32_t bf[2][4];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const uint32_t* p = xw + (32 * s + 16 * h + q) * (K::XSTRIDE / 4) + 8 * u + g;
                    const uint32_t r0 = p[0], r1 = p[4 * (K::XSTRIDE / 4)], r2 = p[8 * (K::XSTRIDE / 4)],
                                   r3 = p[12 * (K::XSTRIDE / 4)];
                    const uint32_t t0 = __byte_perm(r0, r1, 0x5140), t1 = __byte_perm(r2, r3, 0x5140);
                    const uint32_t t2 = __byte_perm(r0, r1, 0x7362), t3 = __byte_perm(r2, r3, 0x7362);
                    bf[h][0] = __byte_perm(t0, t1, 0x5410);
                    bf[h][1] = __byte_perm(t0, t1, 0x7632);
                    bf[h][2] = __byte_perm(t2, t3, 0x5410);
                    bf[h][3] = __byte_perm(t2, t3, 0x7632);
                }
#pragma unroll
                for (int j = 0; j < 4; j++) mma_k32(acc[j], afrag[s], bf[0][j], bf[1][j]);
            }
            // requant + pack: this thread owns samples 32u + 8q .. +7 of filters g and g+8
            uint32_t w[2][2];
#pragma unroll
            for (int c = 0; c < 2; c++) {
                w[0][c] = pack4_sat(tdiv_magic(acc[0][c], m1_lo, sh1_lo), tdiv_magic(acc[1][c], m1_lo, sh1_lo),
                                    tdiv_magic(acc[2][c], m1_lo, sh1_lo), tdiv_magic(acc[3][c], m1_lo, sh1_lo));
                w[1][c] = pack4_sat(tdiv_magic(acc[0][2 + c], m1_hi, sh1_hi), tdiv_magic(acc[1][2 + c], m1_hi, sh1_hi),
                                    tdiv_magic(acc[2][2 + c], m1_hi, sh1_hi), tdiv_magic(acc[3][2 + c], m1_hi, sh1_hi));
            }
            if (K::T % 32 != 0 && u == K::UNITS - 1) {   // zero the samples beyond T (they are layer-2 padding)
#pragma unroll
                for (int c = 0; c < 2; c++) {
                    const int t0 = 32 * u + 8 * q + 4 * c;
                    const uint32_t mask = t0 + 3 < K::T ? 0xffffffffu : (t0 >= K::T ? 0u : (0xffffffffu >> (8 * (4 - (K::T - t0)))));
                    w[0][c] &= mask;
                    w[1][c] &= mask;
                }
            }
            *reinterpret_cast<uint2*>(y1 + g * K::Y1S + 32 + 32 * u + 8 * q) = make_uint2(w[0][0], w[0][1]);
            *reinterpret_cast<uint2*>(y1 + (g + 8) * K::Y1S + 32 + 32 * u + 8 * q) = make_uint2(w[1][0], w[1][1]);
        }
        __syncthreads();   // y1 complete; xs[st] consumed; y4 of the previous trial complete

        if (tid == 0) {    // refill this stage with the trial two iterations ahead
            const long long bn = b + 2 * stride;
            if (bn < B) {
                mbar_expect_tx(bar0 + 8 * st, K::TRIAL_BYTES);
                bulk_g2s(smem_u32(xs + st * K::XS_BYTES), x + (size_t)bn * K::TRIAL_BYTES, K::TRIAL_BYTES, bar0 + 8 * st);
            }
        }

        // ---- layer 5 of the previous trial (one warp)
        if (b_prev >= 0 && warp == kWarps - 1) {
            int z[4] = {0, 0, 0, 0};
            const uint32_t* y4w = reinterpret_cast<const uint32_t*>(y4);
            const uint32_t* w5w = reinterpret_cast<const uint32_t*>(net.w5p);
            for (int i = lane; i < 4 * K::T64P; i += 32) {
                const uint32_t a = y4w[i];
#pragma unroll
                for (int n = 0; n < 4; n++) z[n] = __dp4a((int)a, (int)w5w[n * 4 * K::T64P + i], z[n]);
            }
#pragma unroll
            for (int n = 0; n < 4; n++) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) z[n] += __shfl_xor_sync(0xffffffffu, z[n], o);
                z[n] = tdiv_magic(z[n] + net.b5[n], net.m5, net.sh5);
            }
            if (lane == 0)
                *reinterpret_cast<uint32_t*>(logits + (size_t)b_prev * 4) = pack4_sat(z[0], z[1], z[2], z[3]);
        }

        // ---- layers 2 + 3: this warp owns channels warp and warp + 8
#pragma unroll 1
        for (int k = warp; k < 16; k += kWarps) {
            const uint8_t* tp = net.w2sh + k * kToepChanBytes;
            uint32_t bt[4][3][2];
#pragma unroll
            for (int v = 0; v < 4; v++)
#pragma unroll
                for (int s = 0; s < 3; s++)
#pragma unroll
                    for (int h = 0; h < 2; h++)
                        bt[v][s][h] = *reinterpret_cast<const uint32_t*>(tp + toep_base[v] + 32 * s + 16 * h);
            const int thr = net.thr2[k], sig = net.sig2[k], soff = net.soff2[k], mm = net.m2[k], ss = net.sh2[k];
            uint8_t* y2k = y2 + k * K::Y2S;
#pragma unroll 1
            for (int mt = 0; mt < K::MT; mt++) {
                int row = 16 * mt + (lane & 7) + 8 * ((lane >> 3) & 1);
                row = row < K::MROWS ? row : K::MROWS - 1;
                const uint32_t abase = smem_u32(y1 + k * K::Y1S + 32 * row + 16 * (lane >> 4));
                int acc[4][4];
#pragma unroll
                for (int v = 0; v < 4; v++) acc[v][0] = acc[v][1] = acc[v][2] = acc[v][3] = 0;
#pragma unroll
                for (int s = 0; s < 3; s++) {
                    uint32_t a[4];
                    ldsm_x4(a, abase + 32 * s);
#pragma unroll
                    for (int v = 0; v < 4; v++) mma_k32(acc[v], a, bt[v][s][0], bt[v][s][1]);
                }
#pragma unroll
                for (int hf = 0; hf < 2; hf++) {
                    int s = 0;
#pragma unroll
                    for (int v = 0; v < 4; v++) s += max(acc[v][2 * hf], thr) + max(acc[v][2 * hf + 1], thr);
                    const int p = 64 * mt + 32 * hf + lane;     // pooled sample index
                    const int o = clip8(tdiv_magic(s * sig + soff, mm, ss));
                    if (p < K::T8) y2k[8 + p] = (uint8_t)o;
                }
            }
            __syncwarp();
            // layer 3 on the row just produced
            const int r = (lane + 1) & 3;
            uint32_t w3[5];
#pragma unroll
            for (int i = 0; i < 5; i++) w3[i] = net.w3sh[(k * 4 + r) * 5 + i];
            for (int t = lane; t < K::T8; t += 32) {
                const uint32_t* src = reinterpret_cast<const uint32_t*>(y2k) + ((t + 1) >> 2);
                int a = 0;
#pragma unroll
                for (int i = 0; i < 5; i++) a = __dp4a((int)src[i], (int)w3[i], a);
                y3[t * 16 + k] = (uint8_t)clip8(tdiv_magic(a, net.m3, net.sh3));
            }
        }
        __syncthreads();   // y3 complete; y1 / y2 free for the next trial
        b_prev = b;
    }

    // ---- drain: layers 4 + 5 of the last trial
    if (b_prev >= 0) {
        for (int j = warp; j < K::T64; j += kWarps) {
            int acc[4] = {0, 0, 0, 0};
            const uint32_t bfrag = reinterpret_cast<const uint32_t*>(y3)[(8 * j + g) * 4 + q];
            mma_k16(acc, net.w4frag[g * 4 + q], net.w4frag[(g + 8) * 4 + q], bfrag);
            const int tl = net.thr4[g], th = net.thr4[g + 8];
            int s_lo = max(acc[0], tl) + max(acc[1], tl);
            int s_hi = max(acc[2], th) + max(acc[3], th);
            s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 1);
            s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 1);
            s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 2);
            s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 2);
            if (q == 0) {
                const int xl = s_lo * net.sig4[g] + net.soff4[g];
                const int xh = s_hi * net.sig4[g + 8] + net.soff4[g + 8];
                y4[g * K::T64P + j] = (uint8_t)clip8(tdiv_magic(xl, net.m4[g], net.sh4[g]));
                y4[(g + 8) * K::T64P + j] = (uint8_t)clip8(tdiv_magic(xh, net.m4[g + 8], net.sh4[g + 8]));
            }
        }
        __syncthreads();
        if (warp == 0) {
            int z[4] = {0, 0, 0, 0};
            const uint32_t* y4w = reinterpret_cast<const uint32_t*>(y4);
            const uint32_t* w5w = reinterpret_cast<const uint32_t*>(net.w5p);
            for (int i = lane; i < 4 * K::T64P; i += 32) {
                const uint32_t a = y4w[i];
#pragma unroll
                for (int n = 0; n < 4; n++) z[n] = __dp4a((int)a, (int)w5w[n * 4 * K::T64P + i], z[n]);
            }
#pragma unroll
            for (int n = 0; n < 4; n++) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) z[n] += __shfl_xor_sync(0xffffffffu, z[n], o);
                z[n] = tdiv_magic(z[n] + net.b5[n], net.m5, net.sh5);
            }
            if (lane == 0)
                *reinterpret_cast<uint32_t*>(logits + (size_t)b_prev * 4) = pack4_sat(z[0], z[1], z[2], z[3]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
template <int C, int T>
static cudaError_t launch_instance(const FusedNet* net_dev, const int8_t* x, long long B, int8_t* logits,
                                   int num_sms, int ctas_per_sm, cudaStream_t stream) {
    using K = Cfg<C, T>;
    auto kern = fused_forward_kernel<K>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
    if (e != cudaSuccess) return e;
    if (ctas_per_sm <= 0) {
        int occ = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, K::SMEM);
        if (e != cudaSuccess) return e;
        ctas_per_sm = occ > 0 ? occ : 1;
    }
    long long grid = (long long)num_sms * ctas_per_sm;
    if (grid > B) grid = B;
    kern<<<(unsigned)grid, kThreads, K::SMEM, stream>>>(net_dev, x, B, logits);
    return cudaGetLastError();
}

#define MIBMI_FUSED_SHAPES(X) X(64, 480) X(38, 480) X(24, 480) X(19, 480) X(16, 480) X(8, 480)

bool fused_has_instance(int C, int T) {
#define X(c, t) if (C == c && T == t) return true;
    MIBMI_FUSED_SHAPES(X)
#undef X
    return false;
}

cudaError_t launch_fused(const FusedNet* net_dev, int C, int T, const int8_t* x, long long B, int8_t* logits,
                         int num_sms, int ctas_per_sm, cudaStream_t stream) {
#define X(c, t) if (C == c && T == t) return launch_instance<c, t>(net_dev, x, B, logits, num_sms, ctas_per_sm, stream);
    MIBMI_FUSED_SHAPES(X)
#undef X
    return cudaErrorInvalidValue;
}

}  // namespace mibmi
