// EEGNet / Q-EEGNet layer order: temporal conv (1, Nf) over every EEG channel FIRST, then the depthwise
// spatial conv -- QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:100-106 (modules), :137-168 (forward).  The reference
// tree holds no integer model of this topology; the integer semantics are the edgeEEGNet ones (golden_model.py:153-379:
// truncating division, clip to [-128, 127], REORDER_BN pooling) applied to this layer order -- oracle/eegnet_np.py.
//
//   L1  v[f][c][t] = sum_j xpad[c][t + j] w1[f][j]   (Nf/2 - 1 zeros | x | Nf/2 zeros);  y1 = clip8(tdiv(v + off1[f], fac1[f]))
//   L2  a[k][t] = sum_c y1[k / D][c][t] w2[k][c];  s = sum_8 max(a, -(off2[k] >> 3));  y2 = clip8(tdiv(s + off2[k], fac2[k]))
//   L3..L5 as in mibmi_generic.cu.
//
// One CTA per trial.  Layers 1 and 2 are fused so that the F1 x C x T intermediate (198 kB at 22 x 1125) never exists: a thread
// owns 4 consecutive samples of 4 temporal filters, walks over the EEG channels, computes the Nf-tap convolution with dp4a on
// ALIGNED words of the padded input row against 4 byte-shifted copies of the taps (output t0 + s = sum_m xpad[t0 + m] W_s[m],
// W_s[m] = w[m - s]), requantises with the hardware division (exact for any factor) and accumulates the spatial conv in
// registers.  CUDA-core (dp4a) kernel: 4 x (Nf + 4) / 4 dp4a per output quadruple and channel.
#include <atomic>
#include <new>
#include <string.h>

#include "../../include/mibmi.h"
#include "mibmi_common.cuh"

namespace mibmi {

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
};

__device__ __forceinline__ int eclip8(int v) { return min(max(v, -128), 127); }

template <int D, bool MAGIC>
__global__ void __launch_bounds__(384)
eegnet_forward_kernel(const EegNet* __restrict__ net, const int8_t* __restrict__ x, long long B, int8_t* __restrict__ logits,
                      int8_t* __restrict__ tap2, int8_t* __restrict__ tap3, int8_t* __restrict__ tap4) {
    extern __shared__ __align__(16) int8_t smem[];
    const int C = net->C, T = net->T, F1 = net->F1, K2 = net->K2, F2 = net->F2, N = net->N, Nf = net->Nf;
    const int T8 = net->T8, T64 = net->T64, XP = net->XP, G = net->G, NW = net->NW;
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
        for (int n = tid; n < N; n += nth) {
            const int8_t* w = net->w5 + (size_t)n * F2 * T64;
            int z = net->b5[n];
            for (int i = 0; i < F2 * T64; i++) z += (int)y4[i] * (int)w[i];
            logits[(size_t)b * N + n] = (int8_t)eclip8(z / net->fac5);
        }
        __syncthreads();
    }
}

}  // namespace mibmi

using namespace mibmi;

struct mibmi_eegnet {
    int device = 0, num_sms = 0;
    int C = 0, T = 0, K2 = 0, F2 = 0, N = 0, D = 0, threads = 256, magic1 = 0;
    size_t smem = 0;
    EegNet* net_dev = nullptr;
    int8_t* w5_dev = nullptr;
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
        if (fa < 4 || fa > 0x7fffffffLL || 128 * sum + oa >= (1LL << 29)) { h->magic1 = 0; continue; }
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
    if (e == cudaSuccess) e = cudaMalloc(&n->net_dev, sizeof(EegNet));
    if (e == cudaSuccess) e = cudaMemcpy(n->net_dev, h, sizeof(EegNet), cudaMemcpyHostToDevice);
    delete h;
    if (e != cudaSuccess) {
        if (n->w5_dev) cudaFree(n->w5_dev);
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
    auto kern = n->D == 1 ? (n->magic1 ? eegnet_forward_kernel<1, true> : eegnet_forward_kernel<1, false>)
                          : (n->magic1 ? eegnet_forward_kernel<2, true> : eegnet_forward_kernel<2, false>);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)n->smem);
    if (e == cudaSuccess) {
        long long grid = (long long)n->num_sms * (n->smem > 100 * 1024 ? 1 : 2);
        if (grid > B) grid = B;
        kern<<<(unsigned)grid, n->threads, n->smem, (cudaStream_t)stream>>>(n->net_dev, x_dev, B, logits_dev, y2, y3, y4);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        n->last_cuda.store((int)e);
        return MIBMI_ERR_CUDA;
    }
    n->launches.fetch_add(1);
    return MIBMI_OK;
}

int mibmi_eegnet_forward_int8(mibmi_eegnet_t* n, const int8_t* x_dev, int64_t B, int8_t* logits_dev, void* stream) {
    return mibmi_eegnet_forward_int8_taps(n, x_dev, B, logits_dev, nullptr, nullptr, nullptr, stream);
}

int64_t mibmi_eegnet_launch_count(const mibmi_eegnet_t* n) { return n ? n->launches.load() : 0; }

}  // extern "C"
