// mibmi_reorder.cu -- Wolf order [n][T][C_ALIGN] -> module order [n][C][T] for rows that no tensor map can describe.
//
// The board receives a trial time-major with the channels padded to a multiple of 4 (edge-eegnet_wolf/data/gen_input_header.py:75-76,
// consumed by net_layer1, src/cl/net/layer1.c:121).  The tcgen05 kernels take that order directly when a row is 16 or 64 bytes
// (mibmi_tc.cu, WOLF); rows of 8 / 20 / 24 / 40 bytes (C = 8, 19, 22 | 24, 38) are transposed here, chunk by chunk into a scratch
// buffer small enough to stay in the 126 MB L2, and the fused module-order kernel runs on the chunk (mibmi_api.cu).
//
// One CTA per trial at a time, several CTAs per SM:
//   1. the trial (T * C_ALIGN bytes) is copied into shared memory with cp.async (16 / 8 / 4 bytes as the trial pitch allows), all of
//      it in flight at once, into one of two buffers: the next trial travels while this one is transposed;
//   2. a thread takes 4 samples x 4 channels as four words (lanes run over the channel groups first: at most 2-way bank
//      conflicts for 5, 6 and 10 words per row), transposes them with 8 PRMTs and writes four time-contiguous words into an image
//      of the output whose rows start on words; the row pitch is an odd number of words, so the four rows of a thread and the
//      rows of its neighbours fall into different banks (pitch T / 4 = 120 words put all of them into one: measured 74 % of the
//      shared-memory wavefronts were conflicts, r1g);
//   3. the image leaves with aligned, coalesced 32-bit stores, a warp per row: a global word is an unaligned word of an image row
//      (two LDS + funnel shift, the shift is 0 when T % 4 == 0) or straddles two rows (bytes, a separate pass); the up to 3 head
//      and tail bytes a trial shares with its neighbours' words are stored bytewise.
#include <cuda_runtime.h>
#include <stdint.h>

#include "mibmi_kernels.h"

namespace mibmi {
namespace {

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

constexpr int kReorderThreads = 512;

// CA4 = words per Wolf-order row (C_ALIGN / 4)
template <int CA4>
__global__ void __launch_bounds__(kReorderThreads) wolf_reorder_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ y, long long n,
                                                           int C, int T) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int in_words = T * CA4, CT = C * T, T4 = (T + 3) >> 2;
    constexpr int LANES = kReorderThreads / CA4;               // sample groups in flight (phase 2)
    const size_t in_pitch = ((size_t)in_words * 4 + 4 * CA4 * 4 + 15) & ~(size_t)15;                    // 4 spare input rows
    uint32_t* sin0 = reinterpret_cast<uint32_t*>(smem);
    uint32_t* sin1 = reinterpret_cast<uint32_t*>(smem + in_pitch);
    uint32_t* simg = reinterpret_cast<uint32_t*>(smem + 2 * in_pitch);
    const int TP4 = T4 | 1;                                    // image row pitch in words, odd: rows 4 apart never share a bank
    const int tid = threadIdx.x, nth = blockDim.x, c4 = tid % CA4, tl = tid / CA4;

    // asynchronous copies of the widest size the trial pitch allows; a trial is all in flight at once
    auto fetch = [&](long long trial, uint32_t* dst) {
        const int8_t* gsrc = x + trial * (long long)in_words * 4;
        const uint32_t sdst = (uint32_t)__cvta_generic_to_shared(dst);
        if ((in_words & 3) == 0) {
            for (int i = tid * 16; i < in_words * 4; i += nth * 16)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst + i), "l"(gsrc + i) : "memory");
        } else if ((in_words & 1) == 0) {
            for (int i = tid * 8; i < in_words * 4; i += nth * 8)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sdst + i), "l"(gsrc + i) : "memory");
        } else {
            for (int i = tid * 4; i < in_words * 4; i += nth * 4)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst + i), "l"(gsrc + i) : "memory");
        }
    };

    int buf = 0;
    if (blockIdx.x < n) fetch(blockIdx.x, sin0);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (long long trial = blockIdx.x; trial < n; trial += gridDim.x, buf ^= 1) {
        const uint32_t* sin = buf ? sin1 : sin0;
        if (trial + gridDim.x < n) fetch(trial + gridDim.x, buf ? sin0 : sin1);   // the next trial travels during this one's transposition
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();      // this trial has landed; the image of the previous one has left

        // 4 samples x 4 channels per thread -> four time-contiguous words of the image (row c at c * TP, TP = 4 * ceil(T / 4))
        // (the channel group of a thread is fixed, its sample groups are LANES apart: constant strides, no division in the loop)
        if (tl < LANES) {
            const uint32_t* r = sin + (tl * 4) * CA4 + c4;
            uint32_t* d = simg + (c4 * 4) * TP4 + tl;
            const bool k1 = c4 * 4 + 1 < C, k2 = c4 * 4 + 2 < C, k3 = c4 * 4 + 3 < C;      // c4 * 4 < C always
#pragma unroll 2
            for (int t4 = tl; t4 < T4; t4 += LANES, r += LANES * 4 * CA4, d += LANES) {
                const uint32_t a0 = r[0], a1 = r[CA4], a2 = r[2 * CA4], a3 = r[3 * CA4];   // sample 4 t4 + j, channels 4 c4 .. + 3
                const uint32_t p0 = prmt(a0, a1, 0x5140), p1 = prmt(a2, a3, 0x5140);        // (a0.b0 a1.b0 a0.b1 a1.b1), same for a2 / a3
                const uint32_t q0 = prmt(a0, a1, 0x7362), q1 = prmt(a2, a3, 0x7362);        // bytes 2 and 3
                d[0] = prmt(p0, p1, 0x5410);
                if (k1) d[TP4] = prmt(p0, p1, 0x7632);
                if (k2) d[2 * TP4] = prmt(q0, q1, 0x5410);
                if (k3) d[3 * TP4] = prmt(q0, q1, 0x7632);
            }
        }
        __syncthreads();

        int8_t* g0 = y + trial * (long long)CT;
        {
            // aligned global words; word m holds output bytes i0 = 4 m - ph .. i0 + 3 of this trial (ph = phase of the trial start)
            const int ph = (int)((uintptr_t)g0 & 3);
            int8_t* gbase = g0 - ph;
            const int first = ph ? 1 : 0, last = (ph + CT) >> 2;   // whole words: [first, last)
            uint32_t* gwo = reinterpret_cast<uint32_t*>(gbase);
            const uint8_t* sb = reinterpret_cast<const uint8_t*>(simg);
            // the whole words inside row c are unaligned words of the image row, all at the same phase --
            // two aligned loads and a funnel shift each.  Words that straddle a row end are left to the pass below.
            for (int c = tid >> 5; c < C; c += nth >> 5) {     // a warp per row (a (row, block-of-32-words) split was slower, r1g)
                const int rs = ph + c * T;                     // row start on the word grid of gbase
                const int m0 = (rs + 3) >> 2, m1 = (rs + T) >> 2, t0 = 4 * m0 - rs;   // words [m0, m1), first one starts at sample t0
                const uint32_t* src = simg + c * TP4;
                const int sh = 8 * t0;
                for (int k = tid & 31; k < m1 - m0; k += 32) gwo[m0 + k] = __funnelshift_r(src[k], src[k + 1], sh);
            }
            // the word holding the last byte of row c (c < C - 1) when it also holds bytes of row c + 1
            for (int c = tid; c < C - 1; c += nth) {
                const int e = ph + (c + 1) * T - 1;            // position of the row's last byte in the word grid
                if ((e & 3) != 3) {
                    const int m = e >> 2, i0 = 4 * m - ph;     // i0 >= 0: T >= 4
                    uint32_t v = 0;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int tj = i0 + j - c * T, wrap = tj >= T;
                        v |= (uint32_t)sb[(c + wrap) * TP4 * 4 + tj - (wrap ? T : 0)] << (8 * j);
                    }
                    if (m >= first && m < last) gwo[m] = v;
                }
            }
            if (tid < 4) {                                     // head bytes (global word 0) and tail bytes (global word `last`)
                if (ph && tid >= ph) gbase[tid] = (int8_t)sb[tid - ph];
                const int i = (last << 2) + tid - ph;          // output byte index, in the last row
                if (last >= first && i < CT) gbase[(last << 2) + tid] = (int8_t)sb[(C - 1) * TP4 * 4 + i - (C - 1) * T];
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

size_t reorder_smem(int CA4, int C, int T) {
    const size_t in = ((size_t)T * CA4 * 4 + 4 * CA4 * 4 + 15) & ~(size_t)15;
    return 2 * in + (((size_t)C * (((T + 3) / 4) | 1) * 4 + 8 + 15) & ~(size_t)15);
}

template <int CA4>
cudaError_t launch_one(const int8_t* x, int8_t* y, long long n, int C, int T, int num_sms, cudaStream_t stream) {
    const size_t smem = reorder_smem(CA4, C, T);
    auto kern = wolf_reorder_kernel<CA4>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = (int)((size_t)220 * 1024 / (smem + 1024));
    per_sm = per_sm < 1 ? 1 : per_sm > 2048 / kReorderThreads ? 2048 / kReorderThreads : per_sm;
    long long grid = (long long)num_sms * per_sm;
    if (grid > n) grid = n;
    kern<<<(unsigned)grid, kReorderThreads, smem, stream>>>(x, y, n, C, T);
    return cudaGetLastError();
}

}  // namespace

bool reorder_has_instance(int C, int T) {
    const int ca4 = ((C + 3) & ~3) / 4;
    if (!(ca4 == 2 || ca4 == 5 || ca4 == 6 || ca4 == 10)) return false;
    return reorder_smem(ca4, C, T) <= (size_t)110 * 1024;       // two CTAs per SM at least
}

cudaError_t launch_wolf_reorder(const int8_t* x, int8_t* y, long long n, int C, int T, int num_sms, cudaStream_t stream) {
    if (n <= 0) return cudaSuccess;
    switch (((C + 3) & ~3) / 4) {
        case 2: return launch_one<2>(x, y, n, C, T, num_sms, stream);
        case 5: return launch_one<5>(x, y, n, C, T, num_sms, stream);
        case 6: return launch_one<6>(x, y, n, C, T, num_sms, stream);
        case 10: return launch_one<10>(x, y, n, C, T, num_sms, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace mibmi
