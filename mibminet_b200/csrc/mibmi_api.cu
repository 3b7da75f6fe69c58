// C-ABI of libmibmi.so (include/mibmi.h): validation, parameter packing, dispatch, host staging.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/mibmi.h"
#include "mibmi_common.cuh"
#include "mibmi_kernels.h"

using namespace mibmi;

struct mibmi_net {
    int device = 0, num_sms = 0;
    int C = 0, T = 0, F2 = 0, N = 0, reorder = 1;
    GenericNet* gen_dev = nullptr;
    int8_t* w5_dev = nullptr;
    PipeNet* pipe_dev = nullptr;       // warp-specialised pipeline kernel (MIBMI_PATH_FUSED)
    bool pipe_ok = false, pipe_fast1 = false, stream_fast3 = false, tc_ok = false, tcr_ok = false, tc_wolf_ok = false, plain = false;
    const char* pipe_why = "";
    StreamNet* stream_dev = nullptr;   // sliding-window streaming kernel (mibmi_forward_stream_int8)
    bool stream_ok = false;
    const char* stream_why = "";
    int stream_mode = MIBMI_STREAM_AUTO;
    int force_path = MIBMI_PATH_AUTO;
    int64_t host_chunk = 0;
    std::atomic<int> last_path{0}, last_kernel{0};
    std::atomic<int> last_cuda{0};
    std::atomic<int64_t> launches{0};
    // host entry point staging (guarded by host_mu)
    std::mutex host_mu;
    cudaStream_t hs[2] = {nullptr, nullptr};
    int8_t* hx_dev[2] = {nullptr, nullptr};
    int8_t* hy_dev[2] = {nullptr, nullptr};
    int8_t* hx_pin[2] = {nullptr, nullptr};
    int8_t* hy_pin[2] = {nullptr, nullptr};
    int64_t h_cap_trials = 0;          // capacity of the staging buffers in trials ...
    size_t h_cap_bytes = 0;            // ... and in bytes of input per buffer (the trial size depends on the layout)
    // Wolf-order rows no tensor map describes: transposed chunk by chunk into a scratch that lives for the duration of
    // the call on the caller's stream (cudaMallocAsync / cudaFreeAsync), so concurrent calls on different streams
    // never share it
    bool wolf_reorder_ok = false, tcr_wolf_ok = false;
    int64_t wolf_chunk = 0;
};

namespace {

struct DeviceGuard {
    int prev = -1;
    cudaError_t err;
    explicit DeviceGuard(int dev) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
    }
    ~DeviceGuard() {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
};

inline int fail_cuda(mibmi_net* n, cudaError_t e) {
    if (n) n->last_cuda.store((int)e);
    return MIBMI_ERR_CUDA;
}
inline int align4(int d) { return (d + 3) / 4 * 4; }
inline int64_t iabs64(int64_t v) { return v < 0 ? -v : v; }

// tdiv(x, d) == (mulhi(x, m) >> s) + (x < 0)  for d >= 4 and |x| < 2^29  (proved in DESIGN.md,
// brute-forced in tests/test_host_logic.py through mibmi_debug_magic).
bool magic_for(int64_t d, int* m, int* s) {
    if (d < 4 || d > 0x7fffffff) return false;
    int lg = 0;
    while ((int64_t(1) << (lg + 1)) <= d) lg++;   // floor(log2 d) >= 2
    const int p = 30 + lg;
    const int64_t mm = (int64_t)(((__int128)1 << p) / d) + 1;
    if (mm >= (int64_t(1) << 31)) return false;
    *m = (int)mm;
    *s = p - 32;
    return true;
}

constexpr int64_t kMagicRange = int64_t(1) << 29;

struct Pooled { int thr, sig, soff, m, sh; };

// requant of a pooled layer (L2 / L4, REORDER_BN): x = sum_8 max(v, thr) + off, y = clip(x / fac)
bool pack_pooled(int64_t vmax, int off, int fac, Pooled* o, const char** why) {
    const int thr = -(off >> 3);
    const int64_t smax = 8 * (vmax > iabs64(thr) ? vmax : iabs64(thr));
    if (smax + iabs64(off) >= kMagicRange) { *why = "pooled accumulator outside the magic-division range"; return false; }
    if (!magic_for(iabs64(fac), &o->m, &o->sh)) { *why = "|factor| < 4 in a pooled layer"; return false; }
    o->thr = thr;
    o->sig = fac < 0 ? -1 : 1;
    o->soff = fac < 0 ? -off : off;
    return true;
}

// requant of a pooled layer in the plain variant (REORDER_BN off; layer2.c:139-210, layer4.c:99-140): every accumulator is divided,
// u = tdiv(v + (off >> 3), fac >> 3) -- the kernels compute tdiv(sig v + soff, |fac >> 3|) -- then ReLU, sum of 8, >> 3
bool pack_pooled_plain(int64_t vmax, int off, int fac, Pooled* o, const char** why) {
    const int d = fac >> 3, oo = off >> 3;                      // C arithmetic shifts, as in the reference
    if (vmax + iabs64(oo) >= kMagicRange) { *why = "pooled accumulator outside the magic-division range"; return false; }
    if (!magic_for(iabs64(d), &o->m, &o->sh)) { *why = "|factor >> 3| < 4 in a pooled layer (plain variant)"; return false; }
    o->thr = 0;
    o->sig = d < 0 ? -1 : 1;
    o->soff = d < 0 ? -oo : oo;
    return true;
}

// magic for the "no shift" form q = mulhi(x, m) + (x < 0), m = floor(2^32 / d) + 1.  With e = m d - 2^32
// in (0, d]: x m / 2^32 = x / d + x e / (d 2^32), so the form is exact while |x| e < 2^32 (the error term stays below
// 1 / d).  Beyond |x| >= 130 d the true quotient is >= 130 in magnitude and the approximation errs only away from zero
// (x > 0: floor(x/d + err) >= floor(x/d) >= 130; x < 0: floor(x/d - err) + 1 <= x/d + 1 <= -129), so both saturate to
// the same clip8 value: exactness is needed on [-130 d, 130 d] only.  Tests: tests/test_host_logic.py::test_fast1_*.
bool magic32_for(int64_t d, int64_t xmax, int* m) {
    if (d < 4 || d > 0x7fffffff) return false;
    const int64_t mm = (int64_t(1) << 32) / d + 1;
    const int64_t e = mm * d - (int64_t(1) << 32);
    const int64_t lim = xmax < 130 * d ? xmax : 130 * d;
    if ((__int128)lim * e >= ((__int128)1 << 32)) return false;    // strict: x e == 2^32 with x mod d == d - 1 would round up
    *m = (int)mm;
    return true;
}


bool pack_pipe(const mibmi_params_t* p, PipeNet* f, const char** why) {
    const int C = p->C, T = p->T, F2 = p->F2, CA4 = align4(C);
    const int T64 = (T / 8) / 8, T64A = align4(T64), T64P = (T64 + 1) / 2 * 2;
    memset(f, 0, sizeof(*f));
    if (F2 != 16 || p->N != 4) { *why = "fused kernel needs F2 == 16 and N == 4"; return false; }
    const bool plain = !p->reorder_bn;       // tcgen05 kernels only (module order); the pipeline / streaming kernels refuse it
    f->plain = plain ? 1 : 0;
    if (!pipe_has_instance(C, T)) { *why = "no fused instantiation for this (C, T)"; return false; }
    // ---- layer 1: sign of the factor folded into weights and offset
    const int half0 = (C + 1) / 2, half1 = C - half0;      // PCfg::CA / CB
    bool fast1 = true;
    int64_t xmax1[16];
    for (int k = 0; k < 16; k++) {
        const int fac = p->l1_factor[k], sg = fac < 0 ? -1 : 1;
        int64_t sum = 0;
        for (int c = 0; c < C; c++) {
            const int w = p->l1_weight_align[k * CA4 + c];
            if (sg < 0 && w == -128) { *why = "weight -128 cannot be negated (negative layer-1 factor)"; return false; }
            sum += iabs64(w);
        }
        xmax1[k] = 128 * sum + iabs64(p->l1_offset[k]);
        if (xmax1[k] >= kMagicRange) { *why = "layer-1 accumulator outside the magic-division range"; return false; }
        int m;
        if (!magic32_for(iabs64(fac), xmax1[k], &m)) fast1 = false;
        f->off1[k] = sg * p->l1_offset[k];
    }
    for (int k = 0; k < 16; k++) {
        const int64_t d = iabs64(p->l1_factor[k]);
        if (fast1) { magic32_for(d, xmax1[k], &f->m1[k]); f->sh1[k] = 0; }
        else if (!magic_for(d, &f->m1[k], &f->sh1[k])) { *why = "|l1_factor| < 4"; return false; }
    }
    f->fast1 = fast1 ? 1 : 0;
    // W1 is the B operand of layer 1: [block][n-tile][lane] = channels (ldmatrix rows) 4q..4q+3 of the
    // block, filter 8 nt + g
    for (int kb = 0; kb < (C + 15) / 16; kb++)
        for (int nt = 0; nt < 2; nt++)
            for (int lane = 0; lane < 32; lane++) {
                const int g = lane >> 2, q = lane & 3, row = 8 * nt + g;
                const int sg = p->l1_factor[row] < 0 ? -1 : 1;
                uint32_t word = 0;
                for (int i = 0; i < 4; i++) {
                    const int r = 4 * q + i;                       // K position in the block = ldmatrix row
                    const int region = (r >> 2) & 1, idx = 8 * kb + 4 * (r >> 3) + (r & 3);
                    // rows that are not 16-byte aligned (T % 16 != 0) go through the realigning loader, whose
                    // scratch keeps the natural channel order; otherwise the two channel halves interleave
                    const bool natural = (T % 16) != 0;
                    const bool valid = natural ? (16 * kb + r < C) : (idx < (region ? half1 : half0));
                    const int c = natural ? 16 * kb + r : (region ? half0 + idx : idx);
                    const int w = valid ? sg * p->l1_weight_align[row * CA4 + c] : 0;
                    word |= (uint32_t)(uint8_t)(int8_t)w << (8 * i);
                }
                f->w1blk[kb][nt][lane] = word;
            }
    {
        const int kpad = (C + 31) / 32 * 32;
        for (int n = 0; n < 16; n++) {
            const int sg = p->l1_factor[n] < 0 ? -1 : 1;
            for (int c = 0; c < C && c < 64; c++)
                f->w1umma[(n >> 3) * (kpad / 16) * 128 + (c >> 4) * 128 + (n & 7) * 16 + (c & 15)] =
                    (uint8_t)(int8_t)(sg * p->l1_weight_align[n * CA4 + c]);
        }
    }
    {   // Wolf-order super-rows (PipeNet::w1x): only when the rows are not a multiple of 16 bytes and the block structure fits
        int gcd = CA4, b16 = 16;
        while (b16) { const int t = gcd % b16; gcd = b16; b16 = t; }
        const int S = CA4 % 16 ? 16 / gcd : 1, kp = (S * CA4 + 31) / 32 * 32;
        if (S > 1 && 16 * S * kp <= kW1X)
            for (int pp = 0; pp < S; pp++)
                for (int n = 0; n < 16; n++) {
                    const int sg = p->l1_factor[n] < 0 ? -1 : 1, nn = 16 * pp + n;
                    for (int c = 0; c < C; c++) {
                        const int k = CA4 * pp + c;
                        f->w1x[(nn >> 3) * (kp / 16) * 128 + (k >> 4) * 128 + (nn & 7) * 16 + (k & 15)] =
                            (uint8_t)(int8_t)(sg * p->l1_weight_align[n * CA4 + c]);
                    }
                }
    }
    // ---- layer 2: Toeplitz B fragments.  Window-relative sample j = o + 1 + tap (o = output offset in
    // the 32-sample window), K position kk of k-step s <-> j = 32 s + kk
    for (int k = 0; k < 16; k++) {
        const int8_t* w = p->l2_weight_reverse + k * 64;
        int64_t sum = 0;
        for (int j = 0; j < 64; j++) sum += iabs64(w[j]);
        Pooled po;
        if (!(plain ? pack_pooled_plain(128 * sum, p->l2_offset[k], p->l2_factor[k], &po, why)
                    : pack_pooled(128 * sum, p->l2_offset[k], p->l2_factor[k], &po, why))) return false;
        f->thr2[k] = po.thr; f->sig2[k] = po.sig; f->soff2[k] = po.soff; f->m2[k] = po.m; f->sh2[k] = po.sh;
        for (int n = 0; n < 16; n++)
            for (int kk = 0; kk < 96; kk++) {
                const int tap = kk - n - 1;
                f->w2umma[k * 1536 + (n >> 3) * 768 + (kk >> 4) * 128 + (n & 7) * 16 + (kk & 15)] =
                    (tap >= 0 && tap < 64) ? (uint8_t)w[tap] : 0;
            }
        for (int n = 0; n < 32; n++)        // N = 32, reversed outputs, overlapping core matrices (equal values where they overlap)
            for (int kk = 0; kk < 96; kk++) {
                const int tap = kk + n - 32;
                f->w2umma32[k * kW2U32 + (n >> 3) * 128 + (kk >> 4) * 256 + (n & 7) * 16 + (kk & 15)] =
                    (tap >= 0 && tap < 64) ? (uint8_t)w[tap] : 0;
            }
        for (int v = 0; v < 4; v++)
            for (int s = 0; s < 3; s++)
                for (int h = 0; h < 2; h++)
                    for (int lane = 0; lane < 32; lane++) {
                        const int g = lane >> 2, q = lane & 3;
                        const int o = 8 * (g >> 1) + 2 * v + (g & 1);
                        uint32_t word = 0;
                        for (int i = 0; i < 4; i++) {
                            const int j = 32 * s + 16 * h + 4 * q + i;
                            const int tap = j - o - 1;
                            const int val = (tap >= 0 && tap < 64) ? w[tap] : 0;
                            word |= (uint32_t)(uint8_t)(int8_t)val << (8 * i);
                        }
                        f->w2frag[k][v * 6 + s * 2 + h][lane] = word;
                    }
    }
    // ---- layer 3 (exported array is flipped: torch-order tap j = l3_weight[15 - j]); kk = o + tap + 1
    {
        const int sg = p->l3_factor < 0 ? -1 : 1;
        if (!magic_for(iabs64(p->l3_factor), &f->m3, &f->sh3)) { *why = "|l3_factor| < 4"; return false; }
        {   // "no shift" form (magic32_for above) when it is exact up to saturation: every kernel clips the layer-3 quotient to int8
            int64_t xmax3 = 0;
            for (int k = 0; k < 16; k++) {
                int64_t sum = 0;
                for (int j = 0; j < 16; j++) sum += iabs64(p->l3_weight[k * 16 + j]);
                if (128 * sum > xmax3) xmax3 = 128 * sum;
            }
            int m32;
            if (magic32_for(iabs64(p->l3_factor), xmax3, &m32)) { f->m3 = m32; f->sh3 = 0; }
        }
        for (int k = 0; k < 16; k++)
            for (int v = 0; v < 2; v++)
                for (int h = 0; h < 2; h++)
                    for (int lane = 0; lane < 32; lane++) {
                        const int g = lane >> 2, q = lane & 3;
                        const int o = 8 * v + g;
                        uint32_t word = 0;
                        for (int i = 0; i < 4; i++) {
                            const int tap = 16 * h + 4 * q + i - o - 1;
                            int val = (tap >= 0 && tap < 16) ? p->l3_weight[k * 16 + 15 - tap] : 0;
                            if (sg < 0 && val == -128) { *why = "weight -128 cannot be negated (negative layer-3 factor)"; return false; }
                            word |= (uint32_t)(uint8_t)(int8_t)(sg * val) << (8 * i);
                        }
                        f->w3frag[k][v * 2 + h][lane] = word;
                    }
    }
    // ---- layer 4
    for (int k = 0; k < 16; k++) {
        const int8_t* w = p->l4_weight + k * 16;
        int64_t sum = 0;
        for (int i = 0; i < 16; i++) sum += iabs64(w[i]);
        Pooled po;
        if (!(plain ? pack_pooled_plain(128 * sum, p->l4_offset[k], p->l4_factor[k], &po, why)
                    : pack_pooled(128 * sum, p->l4_offset[k], p->l4_factor[k], &po, why))) return false;
        f->thr4[k] = po.thr; f->sig4[k] = po.sig; f->soff4[k] = po.soff; f->m4[k] = po.m; f->sh4[k] = po.sh;
    }
    for (int hf = 0; hf < 2; hf++)
        for (int lane = 0; lane < 32; lane++) {
            const int g = lane >> 2, q = lane & 3;
            uint32_t word = 0;
            for (int i = 0; i < 4; i++) word |= (uint32_t)(uint8_t)p->l4_weight[(g + 8 * hf) * 16 + 4 * q + i] << (8 * i);
            f->w4frag[hf][lane] = word;
        }
    // ---- layer 5: A fragments, K index kk = ch * T64P + t
    {
        const int sg = p->l5_factor < 0 ? -1 : 1;
        if (!magic_for(iabs64(p->l5_factor), &f->m5, &f->sh5)) { *why = "|l5_factor| < 4"; return false; }
        if (T64P / 2 > kPipeMaxKst5) { *why = "T too large for the fused classifier"; return false; }
        for (int n = 0; n < 4; n++) {
            int64_t sum = 0;
            for (int k = 0; k < 16; k++)
                for (int t = 0; t < T64; t++) {
                    const int w = p->l5_weight[(n * 16 + k) * T64A + t];
                    if (sg < 0 && w == -128) { *why = "weight -128 cannot be negated (negative layer-5 factor)"; return false; }
                    sum += iabs64(w);
                }
            if (128 * sum + 128 >= kMagicRange) { *why = "layer-5 accumulator outside the magic-division range"; return false; }
            f->b5[n] = sg * p->l5_bias[n];
        }
        for (int ks = 0; ks < T64P / 2; ks++)
            for (int h = 0; h < 2; h++)
                for (int lane = 0; lane < 32; lane++) {
                    const int g = lane >> 2, q = lane & 3;
                    uint32_t word = 0;
                    if (g < 4)
                        for (int i = 0; i < 4; i++) {
                            const int kk = 32 * ks + 16 * h + 4 * q + i, ch = kk / T64P, t = kk % T64P;
                            const int w = (t < T64) ? sg * p->l5_weight[(g * 16 + ch) * T64A + t] : 0;
                            word |= (uint32_t)(uint8_t)(int8_t)w << (8 * i);
                        }
                    f->w5frag[ks][h][lane] = word;
                }
    }
    return true;
}

// Streaming kernel: shared parts come from the pipeline packing; window-specific zero padding becomes masked
// Toeplitz fragments (see StreamNet in mibmi_common.cuh).
bool pack_stream(const mibmi_params_t* p, const PipeNet* pn, StreamNet* f, const char** why) {
    const int C = p->C, T = p->T, CA4 = align4(C), T8 = T / 8;
    memset(f, 0, sizeof(*f));
    if (!stream_has_instance(C, T)) { *why = "no streaming instantiation for this (C, T)"; return false; }
    memcpy(f->off1, pn->off1, sizeof(f->off1)); memcpy(f->m1, pn->m1, sizeof(f->m1)); memcpy(f->sh1, pn->sh1, sizeof(f->sh1));
    f->fast1 = pn->fast1;
    memcpy(f->w2frag, pn->w2frag, sizeof(f->w2frag));
    memcpy(f->thr2, pn->thr2, sizeof(f->thr2)); memcpy(f->sig2, pn->sig2, sizeof(f->sig2)); memcpy(f->soff2, pn->soff2, sizeof(f->soff2));
    memcpy(f->m2, pn->m2, sizeof(f->m2)); memcpy(f->sh2, pn->sh2, sizeof(f->sh2));
    f->m3 = pn->m3; f->sh3 = pn->sh3;
    memcpy(f->w4frag, pn->w4frag, sizeof(f->w4frag));
    memcpy(f->thr4, pn->thr4, sizeof(f->thr4)); memcpy(f->sig4, pn->sig4, sizeof(f->sig4)); memcpy(f->soff4, pn->soff4, sizeof(f->soff4));
    memcpy(f->m4, pn->m4, sizeof(f->m4)); memcpy(f->sh4, pn->sh4, sizeof(f->sh4));
    memcpy(f->w5frag, pn->w5frag, sizeof(f->w5frag));
    memcpy(f->b5, pn->b5, sizeof(f->b5)); f->m5 = pn->m5; f->sh5 = pn->sh5;
    // layer 1, natural channel order: K position r of block kb <-> channel 16 kb + r
    for (int kb = 0; kb < (C + 15) / 16; kb++)
        for (int nt = 0; nt < 2; nt++)
            for (int lane = 0; lane < 32; lane++) {
                const int g = lane >> 2, q = lane & 3, row = 8 * nt + g;
                const int sg = p->l1_factor[row] < 0 ? -1 : 1;
                uint32_t word = 0;
                for (int i = 0; i < 4; i++) {
                    const int c = 16 * kb + 4 * q + i;
                    const int w = c < C ? sg * p->l1_weight_align[row * CA4 + c] : 0;
                    word |= (uint32_t)(uint8_t)(int8_t)w << (8 * i);
                }
                f->w1nat[kb][nt][lane] = word;
            }
    // layer 2 edges.  Left: output o of the window from its samples ii = 0..63, tap = ii - o + 31 (31 zeros of
    // left padding, gen_net_header.py:145).  Right: output T - 32 + o from samples T - 64 + ii, tap = ii - o - 1.
    for (int k = 0; k < 16; k++) {
        const int8_t* w = p->l2_weight_reverse + k * 64;
        for (int v = 0; v < 4; v++)
            for (int s = 0; s < 2; s++)
                for (int h = 0; h < 2; h++)
                    for (int lane = 0; lane < 32; lane++) {
                        const int g = lane >> 2, q = lane & 3;
                        const int o = 8 * (g >> 1) + 2 * v + (g & 1);
                        uint32_t wl = 0, wr = 0;
                        for (int i = 0; i < 4; i++) {
                            const int ii = 32 * s + 16 * h + 4 * q + i;
                            const int tl = ii - o + 31, tr = ii - o - 1;
                            wl |= (uint32_t)(uint8_t)((tl >= 0 && tl < 64) ? w[tl] : 0) << (8 * i);
                            wr |= (uint32_t)(uint8_t)((tr >= 0 && tr < 64) ? w[tr] : 0) << (8 * i);
                        }
                        f->w2L[k][v * 4 + s * 2 + h][lane] = wl;
                        f->w2R[k][v * 4 + s * 2 + h][lane] = wr;
                    }
    }
    // layer 3, window = MMA row.  Block b covers outputs 16 b .. 16 b + 15 from pooled samples p = 16 b - 8 + kk;
    // torch-order tap jt = p - t + 7 (7 zeros of left padding, gen_net_header.py:171), exported array flipped.
    // Samples p < 4 and p > T8 - 5 are window specific: masked here, supplied through w3e.
    {
        const int sg = p->l3_factor < 0 ? -1 : 1;
        const int nb3 = (T8 + 15) / 16;
        for (int k = 0; k < 16; k++) {
            for (int cls = 0; cls < 3; cls++)
                for (int e = 0; e < 2; e++)
                    for (int h = 0; h < 2; h++)
                        for (int lane = 0; lane < 32; lane++) {
                            const int c = lane >> 2, q = lane & 3;
                            const int b = cls == 0 ? 0 : (cls == 2 ? nb3 - 1 : 1);
                            uint32_t word = 0;
                            for (int i = 0; i < 4; i++) {
                                const int kk = 16 * h + 4 * q + i, pp = 16 * b - 8 + kk;
                                const int jt = kk - 8 * e - c - 1;
                                const bool interior = cls == 1 || (pp >= 4 && pp <= T8 - 5);
                                const int val = (jt >= 0 && jt < 16 && interior) ? sg * p->l3_weight[k * 16 + 15 - jt] : 0;
                                word |= (uint32_t)(uint8_t)(int8_t)val << (8 * i);
                            }
                            f->w3s[k][cls * 4 + e * 2 + h][lane] = word;
                        }
            for (int idx = 0; idx < 4; idx++)
                for (int lane = 0; lane < 32; lane++) {
                    const int c = lane >> 2, q = lane & 3;
                    const int n = idx < 2 ? idx : 2 * (nb3 - 1) + (idx - 2);
                    uint32_t word = 0;
                    for (int i = 0; i < 4; i++) {
                        const int kk = 4 * q + i;
                        const int pe = kk < 4 ? kk : (kk < 8 ? T8 - 8 + kk : -1000);
                        const int jt = pe - (8 * n + c) + 7;
                        const int val = (jt >= 0 && jt < 16) ? sg * p->l3_weight[k * 16 + 15 - jt] : 0;
                        word |= (uint32_t)(uint8_t)(int8_t)val << (8 * i);
                    }
                    f->w3e[k][idx][lane] = word;
                }
        }
    }
    return true;
}

int validate(const mibmi_params_t* p) {
    if (!p->l1_factor || !p->l1_offset || !p->l1_weight_align || !p->l2_factor || !p->l2_offset ||
        !p->l2_weight_reverse || !p->l3_weight || !p->l4_factor || !p->l4_offset || !p->l4_weight ||
        !p->l5_bias || !p->l5_weight)
        return MIBMI_ERR_NULL;
    if (p->C < 1 || p->C > kMaxC || p->T < 64 || p->T > (1 << 16)) return MIBMI_ERR_SHAPE;
    if (p->F2 < 4 || p->F2 > kMaxF2 || p->F2 % 4 != 0) return MIBMI_ERR_SHAPE;       // gen_net_header.py:63
    if (p->D != 1 || p->F2 != p->F1 * p->D) return MIBMI_ERR_SHAPE;                   // golden_model.py:83,169
    if (p->Nf != kNf) return MIBMI_ERR_SHAPE;                                         // golden_model.py:216-219
    if (p->N < 1 || p->N > kMaxN) return MIBMI_ERR_SHAPE;
    if (generic_smem_bytes(p->T, p->F2) > 220 * 1024) return MIBMI_ERR_SHAPE;
    if (p->l3_factor == 0 || p->l5_factor == 0) return MIBMI_ERR_PARAM;
    for (int k = 0; k < p->F2; k++) {
        if (p->l1_factor[k] == 0 || p->l2_factor[k] == 0 || p->l4_factor[k] == 0) return MIBMI_ERR_PARAM;
        if (!p->reorder_bn && ((p->l2_factor[k] >> 3) == 0 || (p->l4_factor[k] >> 3) == 0)) return MIBMI_ERR_PARAM;
    }
    return MIBMI_OK;
}

void free_host_staging(mibmi_net* n) {
    for (int i = 0; i < 2; i++) {
        if (n->hx_dev[i]) cudaFree(n->hx_dev[i]);
        if (n->hy_dev[i]) cudaFree(n->hy_dev[i]);
        if (n->hx_pin[i]) cudaFreeHost(n->hx_pin[i]);
        if (n->hy_pin[i]) cudaFreeHost(n->hy_pin[i]);
        n->hx_dev[i] = n->hy_dev[i] = n->hx_pin[i] = n->hy_pin[i] = nullptr;
    }
    n->h_cap_trials = 0;
    n->h_cap_bytes = 0;
}

size_t trial_bytes(const mibmi_net* n, int layout) {
    return layout == MIBMI_LAYOUT_BCT ? (size_t)n->C * n->T : (size_t)n->T * align4(n->C);
}

// Names mibmi_last_kernel reports; the index is stored by whoever launches (never inferred from the path afterwards).
enum { K_NONE = 0, K_TC = 1, K_PIPE = 2, K_GENERIC = 3, K_STREAM = 4, K_TCR = 5 };

// The fused module-order kernels, best first.  *which = the kernel that was launched.
cudaError_t launch_fused_bct(mibmi_net* net, int path, const int8_t* x, int64_t B, int8_t* logits, const PeerOut& peers,
                             int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, cudaStream_t stream, int* which) {
    if (path == MIBMI_PATH_FUSED && net->tc_ok) {
        *which = K_TC;
        return launch_tc(net->pipe_dev, net->pipe_fast1 && net->stream_fast3, net->C, net->T, x, B, logits, peers, t1, t2, t3, t4, net->num_sms, stream,
                         net->plain);
    }
    if (path == MIBMI_PATH_FUSED && net->tcr_ok) {
        *which = K_TCR;
        return launch_tcr(net->pipe_dev, net->pipe_fast1 && net->stream_fast3, net->C, net->T, x, B, MIBMI_LAYOUT_BCT, logits, peers, t1, t2, t3, t4,
                          net->num_sms, stream, net->plain);
    }
    if (net->plain) return cudaErrorNotSupported;       // the mma.sync pipeline kernel implements REORDER_BN only
    *which = K_PIPE;
    return launch_pipe(net->pipe_dev, net->pipe_fast1, net->C, net->T, x, B, logits, peers, t1, t2, t3, t4, net->num_sms, stream);
}

// Wolf order through a reorder scratch: transpose a chunk, run the fused kernel on it, next chunk.  The scratch belongs to
// this call: allocated and freed in stream order on the caller's stream (no handle-owned state, safe for concurrent calls
// on different streams).
cudaError_t launch_wolf_chunks(mibmi_net* net, int path, const int8_t* x, int64_t B, int8_t* logits, const PeerOut& peers,
                               int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, cudaStream_t stream, int* which, int64_t* launches) {
    const size_t CT = (size_t)net->C * net->T, in_bytes = (size_t)net->T * align4(net->C);
    const int64_t chunk = B < net->wolf_chunk ? B : net->wolf_chunk;
    int8_t* scratch = nullptr;
    cudaError_t e = cudaMallocAsync(&scratch, CT * (size_t)chunk + 16, stream);
    if (e != cudaSuccess) return e;
    const int T8 = net->T / 8, T64 = T8 / 8;
    for (int64_t off = 0; off < B && e == cudaSuccess; off += chunk) {
        const int64_t nb = B - off < chunk ? B - off : chunk;
        e = launch_wolf_reorder(x + off * in_bytes, scratch, nb, net->C, net->T, net->num_sms, stream);
        if (e != cudaSuccess) break;
        PeerOut po = peers;
        for (int i = 0; i < po.n; i++) po.p[i] += off * net->N;
        e = launch_fused_bct(net, path, scratch, nb, logits + off * net->N, po,
                             t1 ? t1 + off * net->F2 * net->T : nullptr, t2 ? t2 + off * net->F2 * T8 : nullptr,
                             t3 ? t3 + off * net->F2 * T8 : nullptr, t4 ? t4 + off * net->F2 * T64 : nullptr, stream, which);
        *launches += 2;
    }
    const cudaError_t ef = cudaFreeAsync(scratch, stream);
    return e != cudaSuccess ? e : ef;
}

// One launch may cover at most this many trials: the tensor-map coordinates of the TMA kernels (b * T + t, b * C + c) are
// 32-bit.  A multiple of 16, so that every sub-batch of a 16-byte-aligned batch is 16-byte aligned again.
constexpr int64_t kMaxLaunchTrials = int64_t(1) << 20;

int forward_impl(mibmi_net* net, const int8_t* x, int64_t B, int layout, int8_t* logits, int8_t* t1, int8_t* t2,
                 int8_t* t3, int8_t* t4, bool taps, cudaStream_t stream, const PeerOut* peers = nullptr) {
    if (!net) return MIBMI_ERR_NULL;
    if (B < 0 || (layout != MIBMI_LAYOUT_BCT && layout != MIBMI_LAYOUT_BTC)) return MIBMI_ERR_PARAM;
    if (B == 0) return MIBMI_OK;   // empty batch: nothing to read or write, pointers may be null
    if (!x || !logits) return MIBMI_ERR_NULL;
    const bool aligned = ((uintptr_t)x & 15) == 0 && (net->N != 4 || ((uintptr_t)logits & 3) == 0);
    int path = net->force_path;
    // tcgen05 kernel, K-major operand straight from the Wolf order (rows of 64 / 16 bytes as tiles, rows of 40 / 24 / 20 / 8 bytes as super-rows)
    const bool wolf_fused = net->tc_wolf_ok && layout == MIBMI_LAYOUT_BTC && net->force_path != MIBMI_PATH_FUSED_PIPE;
    const bool wolf_tcr = net->tcr_wolf_ok && layout == MIBMI_LAYOUT_BTC && net->force_path != MIBMI_PATH_FUSED_PIPE;   // same, rows of 24 bytes expanded in the kernel
    const bool wolf_reorder = net->wolf_reorder_ok && layout == MIBMI_LAYOUT_BTC && !wolf_fused && !wolf_tcr;   // transpose into a scratch first (pipe kernel only)
    const bool pipe_possible = net->pipe_ok && (layout == MIBMI_LAYOUT_BCT || wolf_reorder || wolf_tcr || wolf_fused);
    // the generic kernel reads bytes: it has no alignment requirement and takes what the fused kernels cannot
    if (path == MIBMI_PATH_AUTO) path = (pipe_possible && !taps && aligned) ? MIBMI_PATH_FUSED : MIBMI_PATH_GENERIC;
    if ((path == MIBMI_PATH_FUSED || path == MIBMI_PATH_FUSED_PIPE) && !pipe_possible) return MIBMI_ERR_UNSUPPORTED;
    if ((path == MIBMI_PATH_FUSED || path == MIBMI_PATH_FUSED_PIPE) && !aligned) return MIBMI_ERR_ALIGN;
    DeviceGuard dg(net->device);
    if (dg.err != cudaSuccess) return fail_cuda(net, dg.err);
    const bool fused = path == MIBMI_PATH_FUSED || path == MIBMI_PATH_FUSED_PIPE;
    const size_t tb = trial_bytes(net, layout);
    const int T8 = net->T / 8, T64 = T8 / 8;
    int which = K_NONE;
    int64_t launches = 0;
    for (int64_t off = 0; off < B; off += kMaxLaunchTrials) {
        const int64_t nb = B - off < kMaxLaunchTrials ? B - off : kMaxLaunchTrials;
        const int8_t* xo = x + (size_t)off * tb;
        int8_t* lo = logits + off * net->N;
        int8_t* o1 = t1 ? t1 + off * net->F2 * net->T : nullptr;
        int8_t* o2 = t2 ? t2 + off * net->F2 * T8 : nullptr;
        int8_t* o3 = t3 ? t3 + off * net->F2 * T8 : nullptr;
        int8_t* o4 = t4 ? t4 + off * net->F2 * T64 : nullptr;
        PeerOut po;
        po.n = 0;
        if (peers) {
            po = *peers;
            for (int i = 0; i < po.n; i++) po.p[i] += off * net->N;
        }
        cudaError_t e;
        if (fused && wolf_fused && path == MIBMI_PATH_FUSED) {
            which = K_TC;
            launches++;
            e = launch_tc_wolf(net->pipe_dev, net->pipe_fast1 && net->stream_fast3, net->C, net->T, xo, nb, lo, po, o1, o2, o3, o4, net->num_sms, stream);
        } else if (fused && wolf_tcr) {
            which = K_TCR;
            launches++;
            e = launch_tcr(net->pipe_dev, net->pipe_fast1 && net->stream_fast3, net->C, net->T, xo, nb, MIBMI_LAYOUT_BTC, lo, po, o1, o2, o3, o4,
                           net->num_sms, stream);
        } else if (fused && wolf_reorder) {
            e = launch_wolf_chunks(net, path, xo, nb, lo, po, o1, o2, o3, o4, stream, &which, &launches);
        } else if (fused) {
            launches++;
            e = launch_fused_bct(net, path, xo, nb, lo, po, o1, o2, o3, o4, stream, &which);
        } else {
            which = K_GENERIC;
            launches++;
            e = launch_generic(net->gen_dev, net->T, net->F2, xo, nb, layout, lo, o1, o2, o3, o4, net->num_sms, stream);
            for (int i = 0; i < po.n && e == cudaSuccess; i++)    // no fused gather in this kernel: peer copies on the same stream
                e = cudaMemcpyAsync(po.p[i], lo, (size_t)nb * net->N, cudaMemcpyDefault, stream);
        }
        if (e != cudaSuccess) return fail_cuda(net, e);
    }
    net->last_path.store(path);
    net->last_kernel.store(which);
    net->launches.fetch_add(launches);
    return MIBMI_OK;
}

}  // namespace

extern "C" {

int mibmi_version(void) { return MIBMI_VERSION; }

const char* mibmi_strerror(int status) {
    switch (status) {
        case MIBMI_OK: return "ok";
        case MIBMI_ERR_NULL: return "null pointer argument";
        case MIBMI_ERR_SHAPE: return "unsupported or inconsistent network dimensions";
        case MIBMI_ERR_PARAM: return "invalid parameter value";
        case MIBMI_ERR_ALIGN: return "device pointer not sufficiently aligned";
        case MIBMI_ERR_CUDA: return "CUDA error or no CUDA device";
        case MIBMI_ERR_UNSUPPORTED: return "not available on the forced kernel path";
        case MIBMI_ERR_TOO_LARGE: return "request exceeds the fixed scratch of the allocation-free entry point (MIBMI_NOALLOC_SCRATCH_BYTES)";
        default: return "unknown status";
    }
}

int mibmi_last_cuda_error(const mibmi_net_t* net) { return net ? net->last_cuda.load() : 0; }
int mibmi_last_path(const mibmi_net_t* net) { return net ? net->last_path.load() : 0; }
const char* mibmi_last_kernel(const mibmi_net_t* net) {
    static const char* const names[] = {"", "tc_forward_kernel", "pipe_forward_kernel", "generic_forward_kernel",
                                        "stream_forward_kernel", "tcr_forward_kernel"};      // indexed by K_*
    const int k = net ? net->last_kernel.load() : 0;
    return names[k >= 0 && k <= K_TCR ? k : 0];
}
int64_t mibmi_launch_count(const mibmi_net_t* net) { return net ? net->launches.load() : 0; }

int mibmi_fused_supported(const mibmi_net_t* net, const char** why) {
    if (!net) return 0;
    if (why) *why = net->pipe_ok ? "" : net->pipe_why;
    return net->pipe_ok ? 1 : 0;
}

int mibmi_debug_magic(int32_t d, int32_t* m, int32_t* s) {
    if (!m || !s) return MIBMI_ERR_NULL;
    return magic_for(d, m, s) ? MIBMI_OK : MIBMI_ERR_PARAM;
}

int mibmi_debug_magic32(int32_t d, int64_t xmax, int32_t* m) {
    if (!m) return MIBMI_ERR_NULL;
    if (xmax < 0) return MIBMI_ERR_PARAM;
    return magic32_for(d, xmax, m) ? MIBMI_OK : MIBMI_ERR_UNSUPPORTED;
}

// Diagnostics (no GPU needed): the packed network exactly as the fused kernels read it.
int64_t mibmi_debug_pack_pipe(const mibmi_params_t* p, void* out, int64_t cap, const char** why) {
    if (!p) return MIBMI_ERR_NULL;
    int st = validate(p);
    if (st != MIBMI_OK) return st;
    if (!out) return (int64_t)sizeof(PipeNet);
    if (cap < (int64_t)sizeof(PipeNet)) return MIBMI_ERR_PARAM;
    const char* w = "";
    const bool ok = pack_pipe(p, reinterpret_cast<PipeNet*>(out), &w);
    if (why) *why = w;
    return ok ? (int64_t)sizeof(PipeNet) : MIBMI_ERR_UNSUPPORTED;
}

int mibmi_create(const mibmi_params_t* p, int device, mibmi_net_t** out) {
    if (!p || !out) return MIBMI_ERR_NULL;
    *out = nullptr;
    int st = validate(p);
    if (st != MIBMI_OK) return st;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return MIBMI_ERR_CUDA;
    mibmi_net* n = new (std::nothrow) mibmi_net();
    if (!n) return MIBMI_ERR_CUDA;
    n->device = device;
    n->C = p->C; n->T = p->T; n->F2 = p->F2; n->N = p->N; n->reorder = p->reorder_bn ? 1 : 0;
    DeviceGuard dg(device);
    cudaError_t e = dg.err;
    const int C = p->C, T = p->T, F2 = p->F2, N = p->N, CA = align4(C), T64 = (T / 8) / 8, T64A = align4(T64);
    std::vector<int8_t> w5((size_t)N * F2 * T64);
    GenericNet* g = new (std::nothrow) GenericNet();
    PipeNet* pn = new (std::nothrow) PipeNet();
    StreamNet* sn = new (std::nothrow) StreamNet();
    if (!g || !pn || !sn) { delete g; delete pn; delete sn; delete n; return MIBMI_ERR_CUDA; }
    memset(g, 0, sizeof(*g));
    g->C = C; g->T = T; g->F2 = F2; g->N = N; g->T8 = T / 8; g->T64 = T64; g->reorder = n->reorder;
    for (int k = 0; k < F2; k++) {
        g->fac1[k] = p->l1_factor[k]; g->off1[k] = p->l1_offset[k];
        g->fac2[k] = p->l2_factor[k]; g->off2[k] = p->l2_offset[k];
        g->fac4[k] = p->l4_factor[k]; g->off4[k] = p->l4_offset[k];
        for (int c = 0; c < C; c++) g->w1[k * C + c] = p->l1_weight_align[k * CA + c];
        for (int j = 0; j < kNf; j++) g->w2[k * kNf + j] = p->l2_weight_reverse[k * kNf + j];
        for (int j = 0; j < kL3Taps; j++) g->w3[k * kL3Taps + j] = p->l3_weight[k * kL3Taps + (kL3Taps - 1 - j)];
        for (int i = 0; i < F2; i++) g->w4[k * F2 + i] = p->l4_weight[k * F2 + i];
    }
    g->fac3 = p->l3_factor; g->fac5 = p->l5_factor;
    for (int c = 0; c < N; c++) {
        g->b5[c] = p->l5_bias[c];
        for (int k = 0; k < F2; k++)
            for (int t = 0; t < T64; t++) w5[((size_t)c * F2 + k) * T64 + t] = p->l5_weight[((size_t)c * F2 + k) * T64A + t];
    }
    n->pipe_ok = pack_pipe(p, pn, &n->pipe_why);
    n->pipe_fast1 = pn->fast1 != 0;
    n->stream_fast3 = pn->sh3 == 0;
    n->plain = pn->plain != 0;
    n->tc_ok = n->pipe_ok && tc_has_instance(p->C, p->T);
    n->tcr_ok = n->pipe_ok && tcr_has_instance(p->C, p->T);
    if (n->plain && !n->tc_ok && !n->tcr_ok && n->pipe_ok) {        // the plain variant exists in the tcgen05 kernels only
        n->pipe_ok = false;
        n->pipe_why = "the plain (REORDER_BN off) variant needs a tcgen05 instantiation of this (C, T)";
    }
    n->tc_wolf_ok = n->tc_ok && !n->plain && tc_has_wolf_instance(p->C, p->T);       // plain: module order only
    n->tcr_wolf_ok = n->tcr_ok && !n->plain && tcr_has_wolf_instance(p->C, p->T);
    n->wolf_reorder_ok = n->pipe_ok && !n->plain && reorder_has_instance(p->C, p->T);     // used when the mma.sync pipeline kernel is forced
    if (n->pipe_ok && !n->plain) n->stream_ok = pack_stream(p, pn, sn, &n->stream_why);
    else n->stream_why = n->pipe_ok ? "the streaming kernel implements the REORDER_BN variant only" : n->pipe_why;

    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) n->num_sms = prop.multiProcessorCount;
    {
        // chunk: 16 waves of the persistent grids (6 trials per CTA and wave).  Measured (r1g): the fixed cost per launch of the
        // forward kernels outweighs keeping the chunk L2-resident, so the chunk is sized for launch count, not for the L2.
        const int64_t wave = (int64_t)(n->num_sms > 0 ? n->num_sms : 148) * 6;
        n->wolf_chunk = 16 * wave;
        if (const char* ev = getenv("MIBMI_WOLF_CHUNK"))        // tuning / tests: exact chunk size
            if (atoll(ev) > 0) n->wolf_chunk = atoll(ev);
    }
    if (e == cudaSuccess) e = cudaMalloc(&n->w5_dev, w5.size() ? w5.size() : 1);
    if (e == cudaSuccess) e = cudaMemcpy(n->w5_dev, w5.data(), w5.size(), cudaMemcpyHostToDevice);
    g->w5 = n->w5_dev;
    if (e == cudaSuccess) e = cudaMalloc(&n->gen_dev, sizeof(GenericNet));
    if (e == cudaSuccess) e = cudaMemcpy(n->gen_dev, g, sizeof(GenericNet), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && n->pipe_ok) {
        e = cudaMalloc(&n->pipe_dev, sizeof(PipeNet));
        if (e == cudaSuccess) e = cudaMemcpy(n->pipe_dev, pn, sizeof(PipeNet), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess && n->stream_ok) {
        e = cudaMalloc(&n->stream_dev, sizeof(StreamNet));
        if (e == cudaSuccess) e = cudaMemcpy(n->stream_dev, sn, sizeof(StreamNet), cudaMemcpyHostToDevice);
    }
    delete g;
    delete pn;
    delete sn;
    if (e != cudaSuccess) {
        mibmi_destroy(n);
        return MIBMI_ERR_CUDA;
    }
    *out = n;
    return MIBMI_OK;
}

int mibmi_destroy(mibmi_net_t* net) {
    if (!net) return MIBMI_ERR_NULL;
    {
        DeviceGuard dg(net->device);
        if (net->gen_dev) cudaFree(net->gen_dev);
        if (net->w5_dev) cudaFree(net->w5_dev);
        if (net->pipe_dev) cudaFree(net->pipe_dev);
        if (net->stream_dev) cudaFree(net->stream_dev);
        free_host_staging(net);
        for (int i = 0; i < 2; i++)
            if (net->hs[i]) cudaStreamDestroy(net->hs[i]);
    }
    delete net;
    return MIBMI_OK;
}

int mibmi_set_option(mibmi_net_t* net, int option, int64_t value) {
    if (!net) return MIBMI_ERR_NULL;
    switch (option) {
        case MIBMI_OPT_FORCE_PATH:
            if (value != MIBMI_PATH_AUTO && value != MIBMI_PATH_GENERIC && value != MIBMI_PATH_FUSED && value != MIBMI_PATH_FUSED_PIPE)
                return MIBMI_ERR_PARAM;
            if ((value == MIBMI_PATH_FUSED || value == MIBMI_PATH_FUSED_PIPE) && !net->pipe_ok) return MIBMI_ERR_UNSUPPORTED;
            if (value == MIBMI_PATH_FUSED_PIPE && net->plain) return MIBMI_ERR_UNSUPPORTED;      // REORDER_BN only
            net->force_path = (int)value;
            return MIBMI_OK;
        case MIBMI_OPT_STREAM_MODE:
            if (value < MIBMI_STREAM_AUTO || value > MIBMI_STREAM_GATHER) return MIBMI_ERR_PARAM;
            if (value == MIBMI_STREAM_FUSED && !net->stream_ok) return MIBMI_ERR_UNSUPPORTED;
            net->stream_mode = (int)value;
            return MIBMI_OK;
        case MIBMI_OPT_HOST_CHUNK:
            if (value < 0) return MIBMI_ERR_PARAM;
            net->host_chunk = value;
            return MIBMI_OK;
        default:
            return MIBMI_ERR_PARAM;
    }
}

int mibmi_forward_int8(mibmi_net_t* net, const int8_t* x_dev, int64_t B, int layout, int8_t* logits_dev, void* stream) {
    return forward_impl(net, x_dev, B, layout, logits_dev, nullptr, nullptr, nullptr, nullptr, false, (cudaStream_t)stream);
}

int mibmi_forward_int8_gather(mibmi_net_t* net, const int8_t* x_dev, int64_t B, int layout, int8_t* const* out_dev,
                              int32_t n_out, int64_t row0, void* stream) {
    if (!net || !out_dev) return MIBMI_ERR_NULL;
    if (n_out < 1 || n_out > kMaxPeers + 1 || row0 < 0) return MIBMI_ERR_PARAM;
    PeerOut po;
    po.n = n_out - 1;
    for (int i = 0; i < n_out; i++) {
        if (!out_dev[i]) return MIBMI_ERR_NULL;
        if (i > 0) po.p[i - 1] = out_dev[i] + row0 * net->N;
    }
    return forward_impl(net, x_dev, B, layout, out_dev[0] + row0 * net->N, nullptr, nullptr, nullptr, nullptr, false,
                        (cudaStream_t)stream, &po);
}

int mibmi_forward_int8_taps(mibmi_net_t* net, const int8_t* x_dev, int64_t B, int layout, int8_t* logits_dev,
                            int8_t* y1, int8_t* y2, int8_t* y3, int8_t* y4, void* stream) {
    return forward_impl(net, x_dev, B, layout, logits_dev, y1, y2, y3, y4, true, (cudaStream_t)stream);
}

int mibmi_forward_int8_host(mibmi_net_t* net, const int8_t* x_host, int64_t B, int layout, int8_t* logits_host) {
    if (!net) return MIBMI_ERR_NULL;
    if (B < 0 || (layout != MIBMI_LAYOUT_BCT && layout != MIBMI_LAYOUT_BTC)) return MIBMI_ERR_PARAM;
    if (B == 0) return MIBMI_OK;
    if (!x_host || !logits_host) return MIBMI_ERR_NULL;
    std::lock_guard<std::mutex> lock(net->host_mu);
    DeviceGuard dg(net->device);
    if (dg.err != cudaSuccess) return fail_cuda(net, dg.err);
    const size_t tb = trial_bytes(net, layout);
    int64_t chunk = net->host_chunk > 0 ? net->host_chunk : 8192;
    if (chunk > B) chunk = B;
    cudaError_t e = cudaSuccess;
    // is the caller's memory page-locked?  then copy straight from / to it
    cudaPointerAttributes pa;
    bool x_pinned = cudaPointerGetAttributes(&pa, x_host) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    bool y_pinned = cudaPointerGetAttributes(&pa, logits_host) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();
    // the staging buffers hold `chunk` trials of THIS call's layout: the trial size differs between the layouts
    // (C * T vs T * align4(C)), so capacity is tracked in bytes as well as in trials
    if (net->h_cap_trials < chunk || net->h_cap_bytes < tb * (size_t)chunk) {
        free_host_staging(net);
        for (int i = 0; i < 2 && e == cudaSuccess; i++) {
            if (!net->hs[i]) e = cudaStreamCreateWithFlags(&net->hs[i], cudaStreamNonBlocking);
            if (e == cudaSuccess) e = cudaMalloc(&net->hx_dev[i], tb * chunk);
            if (e == cudaSuccess) e = cudaMalloc(&net->hy_dev[i], (size_t)net->N * chunk);
            if (e == cudaSuccess) e = cudaMallocHost(&net->hx_pin[i], tb * chunk);
            if (e == cudaSuccess) e = cudaMallocHost(&net->hy_pin[i], (size_t)net->N * chunk);
        }
        if (e != cudaSuccess) { free_host_staging(net); return fail_cuda(net, e); }
        net->h_cap_trials = chunk;
        net->h_cap_bytes = tb * (size_t)chunk;
    }
    int status = MIBMI_OK;
    int64_t pending_lo[2] = {-1, -1}, pending_n[2] = {0, 0};
    int64_t done = 0;
    for (int i = 0; done < B || pending_lo[0] >= 0 || pending_lo[1] >= 0; i++) {
        const int s = i & 1;
        if (pending_lo[s] >= 0) {   // retire the chunk that used this slot two steps ago
            e = cudaStreamSynchronize(net->hs[s]);
            if (e != cudaSuccess) { status = fail_cuda(net, e); break; }
            if (!y_pinned) memcpy(logits_host + pending_lo[s] * net->N, net->hy_pin[s], (size_t)pending_n[s] * net->N);
            pending_lo[s] = -1;
        }
        if (done >= B) continue;
        const int64_t nb = B - done < chunk ? B - done : chunk;
        const int8_t* src = x_host + (size_t)done * tb;
        if (!x_pinned) { memcpy(net->hx_pin[s], src, tb * nb); src = net->hx_pin[s]; }
        e = cudaMemcpyAsync(net->hx_dev[s], src, tb * nb, cudaMemcpyHostToDevice, net->hs[s]);
        if (e != cudaSuccess) { status = fail_cuda(net, e); break; }
        status = forward_impl(net, net->hx_dev[s], nb, layout, net->hy_dev[s], nullptr, nullptr, nullptr, nullptr, false, net->hs[s]);
        if (status != MIBMI_OK) break;
        int8_t* dst = y_pinned ? logits_host + (size_t)done * net->N : net->hy_pin[s];
        e = cudaMemcpyAsync(dst, net->hy_dev[s], (size_t)nb * net->N, cudaMemcpyDeviceToHost, net->hs[s]);
        if (e != cudaSuccess) { status = fail_cuda(net, e); break; }
        pending_lo[s] = done;
        pending_n[s] = nb;
        done += nb;
    }
    if (status != MIBMI_OK) {
        cudaStreamSynchronize(net->hs[0]);
        cudaStreamSynchronize(net->hs[1]);
    }
    return status;
}

int64_t mibmi_stream_windows(const mibmi_net_t* net, int64_t L, int32_t hop) {
    if (!net) return MIBMI_ERR_NULL;
    if (hop < 1 || L < net->T) return MIBMI_ERR_PARAM;
    return (L - net->T) / hop + 1;
}

int mibmi_stream_supported(const mibmi_net_t* net, const char** why) {
    if (!net) return 0;
    if (why) *why = net->stream_ok ? "" : net->stream_why;
    return net->stream_ok ? 1 : 0;
}

static int stream_impl(mibmi_net* net, const int8_t* x, int64_t S, int64_t L, int32_t hop, int8_t* logits, int8_t* t2i,
                       int8_t* t2e, int8_t* t3, int8_t* t4, bool taps, cudaStream_t stream) {
    if (!net) return MIBMI_ERR_NULL;
    if (S < 0 || hop < 1 || L < net->T) return MIBMI_ERR_PARAM;
    if (S == 0) return MIBMI_OK;
    if (!x || !logits) return MIBMI_ERR_NULL;
    if (((uintptr_t)x & 15) != 0 || (net->N == 4 && ((uintptr_t)logits & 3) != 0)) return MIBMI_ERR_ALIGN;
    const int64_t W = (L - net->T) / hop + 1;
    const bool fused_possible = net->stream_ok && hop == 16 && L % 16 == 0 && S * net->C < (int64_t(1) << 31) &&
                                L < (int64_t(1) << 31);
    int mode = net->stream_mode;
    if (mode == MIBMI_STREAM_AUTO) mode = fused_possible ? MIBMI_STREAM_FUSED : MIBMI_STREAM_GATHER;
    if (mode == MIBMI_STREAM_FUSED && !fused_possible) return MIBMI_ERR_UNSUPPORTED;
    if (taps && mode != MIBMI_STREAM_FUSED) return MIBMI_ERR_UNSUPPORTED;
    DeviceGuard dg(net->device);
    if (dg.err != cudaSuccess) return fail_cuda(net, dg.err);
    if (mode == MIBMI_STREAM_FUSED) {
        cudaError_t e = launch_stream(net->stream_dev, (net->pipe_fast1 ? 1 : 0) | (net->stream_fast3 ? 2 : 0), net->C, net->T, x, S, L, logits, t2i, t2e, t3, t4,
                                      net->num_sms, stream);
        if (e != cudaSuccess) return fail_cuda(net, e);
        net->last_path.store(MIBMI_PATH_STREAM);
        net->last_kernel.store(K_STREAM);
        net->launches.fetch_add(1);
        return MIBMI_OK;
    }
    // any hop / shape: materialise bounded batches of windows (stream-ordered scratch), then the per-trial path
    const int64_t total = S * W, tb = (int64_t)net->C * net->T;
    int64_t chunk = (int64_t(256) << 20) / tb;
    if (chunk > total) chunk = total;
    int8_t* scratch = nullptr;
    cudaError_t e = cudaMallocAsync(&scratch, (size_t)(chunk * tb), stream);
    if (e != cudaSuccess) return fail_cuda(net, e);
    int status = MIBMI_OK;
    for (int64_t lo = 0; lo < total && status == MIBMI_OK; lo += chunk) {
        const int64_t nw = total - lo < chunk ? total - lo : chunk;
        e = launch_gather_windows(x, L, net->C, net->T, hop, W, lo, nw, scratch, stream);
        if (e != cudaSuccess) { status = fail_cuda(net, e); break; }
        net->launches.fetch_add(1);
        status = forward_impl(net, scratch, nw, MIBMI_LAYOUT_BCT, logits + lo * net->N, nullptr, nullptr, nullptr, nullptr,
                              false, stream);
    }
    cudaFreeAsync(scratch, stream);
    return status;
}

int mibmi_forward_stream_int8(mibmi_net_t* net, const int8_t* x_dev, int64_t S, int64_t L, int32_t hop, int8_t* logits_dev,
                              void* stream) {
    return stream_impl(net, x_dev, S, L, hop, logits_dev, nullptr, nullptr, nullptr, nullptr, false, (cudaStream_t)stream);
}

int mibmi_forward_stream_int8_taps(mibmi_net_t* net, const int8_t* x_dev, int64_t S, int64_t L, int32_t hop,
                                   int8_t* logits_dev, int8_t* y2_stream, int8_t* y2_edges, int8_t* y3, int8_t* y4,
                                   void* stream) {
    return stream_impl(net, x_dev, S, L, hop, logits_dev, y2_stream, y2_edges, y3, y4, true, (cudaStream_t)stream);
}

// ---- allocation-free windowed scoring (no counterpart in the reference; see mibmi.h) ------------------------------------
// The window scratch of the gather path as a file-scope device array.  In HOST code `g_noalloc_scratch` is only a stand-in for
// the symbol, not a device address: its address on each device comes from cudaGetSymbolAddress, asked once per device with that
// device current, and that pointer is what the kernels receive.
static __device__ __align__(256) int8_t g_noalloc_scratch[MIBMI_NOALLOC_SCRATCH_BYTES];
constexpr int kNoallocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static int8_t* g_noalloc_ptr[kNoallocMaxDevices] = {};

static cudaError_t noalloc_scratch_ptr(int device, int8_t** out) {      // the caller has made `device` current
    if (device < 0 || device >= kNoallocMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = static_cast<int8_t*>(p);
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

int mibmi_forward_stream_int8_noalloc(mibmi_net_t* net, const int8_t* x_dev, int64_t S, int64_t L, int32_t hop,
                                      int8_t* logits_dev, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!net) return MIBMI_ERR_NULL;
    if (S < 0 || hop < 1 || L < net->T) return MIBMI_ERR_PARAM;
    if (S == 0) return MIBMI_OK;
    if (!x_dev || !logits_dev) return MIBMI_ERR_NULL;
    if (((uintptr_t)x_dev & 15) != 0 || (net->N == 4 && ((uintptr_t)logits_dev & 3) != 0)) return MIBMI_ERR_ALIGN;
    const int64_t W = (L - net->T) / hop + 1, tb = (int64_t)net->C * net->T;
    if (S > (int64_t(1) << 40) / W || S * W > MIBMI_NOALLOC_SCRATCH_BYTES / tb) return MIBMI_ERR_TOO_LARGE;
    const int64_t total = S * W;
    DeviceGuard dg(net->device);
    if (dg.err != cudaSuccess) return fail_cuda(net, dg.err);
    int8_t* scratch = nullptr;
    cudaError_t e = noalloc_scratch_ptr(net->device, &scratch);
    if (e != cudaSuccess) return fail_cuda(net, e);
    e = launch_gather_windows(x_dev, L, net->C, net->T, hop, W, 0, total, scratch, stream);
    if (e != cudaSuccess) return fail_cuda(net, e);
    net->launches.fetch_add(1);
    return forward_impl(net, scratch, total, MIBMI_LAYOUT_BCT, logits_dev, nullptr, nullptr, nullptr, nullptr, false, stream);
}

int mibmi_quantize_input_f32(const float* x_dev, int64_t n, float abs_max, int8_t* q_dev, void* stream) {
    if (!x_dev || !q_dev) return MIBMI_ERR_NULL;
    if (n < 0 || !(abs_max > 0.0f)) return MIBMI_ERR_PARAM;
    return launch_quantize(x_dev, n, abs_max, q_dev, (cudaStream_t)stream) == cudaSuccess ? MIBMI_OK : MIBMI_ERR_CUDA;
}

int mibmi_dequantize_logits_f32(const int8_t* q_dev, int64_t n, float abs_max, float* y_dev, void* stream) {
    if (!q_dev || !y_dev) return MIBMI_ERR_NULL;
    if (n < 0) return MIBMI_ERR_PARAM;
    return launch_dequantize(q_dev, n, abs_max, y_dev, (cudaStream_t)stream) == cudaSuccess ? MIBMI_OK : MIBMI_ERR_CUDA;
}

}  // extern "C"
