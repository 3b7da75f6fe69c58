/* ORACLE (test infrastructure, not product code): plain-C restatement of the Wolf integer pipeline.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this file.  It exists because the reference C (oracle/_ref) has its network baked in at compile
 * time (net.h / net.c); this port takes the exported arrays at run time, so any synthetic net can be
 * checked at full batch sizes on the GPU box.  Parity status: PINNED -- tests/test_oracle.py checks
 * it against tests/golden/*.npz (outputs of the reference's GoldenModel and of the reference C
 * compiled here, see oracle/gen_golden.py).
 *
 * Follows, layer by layer:
 *   L1  edge-eegnet_wolf/src/cl/net/layer1.c:87-96      golden_model.py:192-196
 *   L2  layer2.c:87-118 (REORDER_BN) / :139-210 (plain)  golden_model.py:241-254
 *   L3  layer3.c:70 (func_conv_scale, conv.c:133-142)    golden_model.py:285-289
 *   L4  layer4.c:99-140 (plain variant: no per-element clip, :115-117)  golden_model.py:330-342
 *   L5  layer5.c:81-89                                   golden_model.py:374-379
 * Integer semantics: '/' truncates toward zero, '>>' is arithmetic, clip = __CLIP_R(x,127) = [-128,127].
 *
 * Parameter arrays are in the exported layout of data/gen_net_header.py:109-224:
 *   l1_weight_align [F2][C_ALIGN], l2_weight_reverse [F2][64] (torch / cross-correlation order),
 *   l3_weight [F2][16] (flipped: true convolution), l4_weight [F2out][F2in],
 *   l5_weight [N][F2][T64_ALIGN], l5_bias [N]; int32 factor/offset per channel for L1, L2, L4.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int C, T, F2, N;
    const int32_t *l1_factor, *l1_offset; const int8_t* l1_weight_align;
    const int32_t *l2_factor, *l2_offset; const int8_t* l2_weight_reverse;
    int32_t l3_factor; const int8_t* l3_weight;
    const int32_t *l4_factor, *l4_offset; const int8_t* l4_weight;
    int32_t l5_factor; const int8_t* l5_bias; const int8_t* l5_weight;
} wolf_port_net;

static inline int32_t clip8(int32_t x) { return x < -128 ? -128 : (x > 127 ? 127 : x); }
static inline int align4(int d) { return (d + 3) / 4 * 4; }

/* one trial; x is [C][T] (module order).  taps (optional) = y1[F2][T] y2[F2][T8] y3[F2][T8] y4[F2][T64] */
static void forward_one(const wolf_port_net* p, const int8_t* x, int reorder_bn, int8_t* out,
                        int8_t* scratch, int8_t* taps) {
    const int C = p->C, T = p->T, F2 = p->F2, N = p->N;
    const int CA = align4(C), T8 = T / 8, T64 = T8 / 8, T64A = align4(T64);
    int8_t* y1 = scratch;                       /* [F2][T + 63], zero padded 31 | T | 32 */
    int8_t* y2 = y1 + (size_t)F2 * (T + 63);    /* [F2][T8 + 15], zero padded 7 | T8 | 8 */
    int8_t* y3 = y2 + (size_t)F2 * (T8 + 15);   /* [F2][T8] */
    int8_t* y4 = y3 + (size_t)F2 * T8;          /* [F2][T64] */
    memset(y1, 0, (size_t)F2 * (T + 63) + (size_t)F2 * (T8 + 15));

    for (int k = 0; k < F2; k++) {              /* ---- layer 1 */
        const int8_t* w = p->l1_weight_align + (size_t)k * CA;
        for (int t = 0; t < T; t++) {
            int32_t a = 0;
            for (int c = 0; c < C; c++) a += (int32_t)x[(size_t)c * T + t] * w[c];
            y1[(size_t)k * (T + 63) + 31 + t] = (int8_t)clip8((a + p->l1_offset[k]) / p->l1_factor[k]);
        }
    }
    for (int k = 0; k < F2; k++) {              /* ---- layer 2 */
        const int8_t* w = p->l2_weight_reverse + (size_t)k * 64;
        const int8_t* r = y1 + (size_t)k * (T + 63);
        const int32_t off = p->l2_offset[k], fac = p->l2_factor[k];
        for (int q = 0; q < T8; q++) {
            int32_t s = 0;
            for (int e = 0; e < 8; e++) {
                int32_t v = 0;
                const int8_t* xw = r + q * 8 + e;
                for (int j = 0; j < 64; j++) v += (int32_t)xw[j] * w[j];
                if (reorder_bn) {
                    int32_t thr = -(off >> 3);
                    s += v > thr ? v : thr;
                } else {
                    int32_t u = clip8((v + (off >> 3)) / (fac >> 3));
                    s += u > 0 ? u : 0;
                }
            }
            y2[(size_t)k * (T8 + 15) + 7 + q] = (int8_t)(reorder_bn ? clip8((s + off) / fac) : clip8(s >> 3));
        }
    }
    for (int k = 0; k < F2; k++) {              /* ---- layer 3: true convolution with the flipped copy */
        const int8_t* w = p->l3_weight + (size_t)k * 16;
        const int8_t* r = y2 + (size_t)k * (T8 + 15);
        for (int t = 0; t < T8; t++) {
            int32_t a = 0;
            for (int j = 0; j < 16; j++) a += (int32_t)r[t + j] * w[15 - j];
            y3[(size_t)k * T8 + t] = (int8_t)clip8(a / p->l3_factor);
        }
    }
    for (int k = 0; k < F2; k++) {              /* ---- layer 4 */
        const int8_t* w = p->l4_weight + (size_t)k * F2;
        const int32_t off = p->l4_offset[k], fac = p->l4_factor[k];
        for (int q = 0; q < T64; q++) {
            int32_t s = 0;
            for (int e = 0; e < 8; e++) {
                int32_t v = 0;
                for (int i = 0; i < F2; i++) v += (int32_t)y3[(size_t)i * T8 + q * 8 + e] * w[i];
                if (reorder_bn) {
                    int32_t thr = -(off >> 3);
                    s += v > thr ? v : thr;
                } else {
                    int32_t u = (v + (off >> 3)) / (fac >> 3);   /* no clip: layer4.c:115-117 */
                    s += u > 0 ? u : 0;
                }
            }
            y4[(size_t)k * T64 + q] = (int8_t)(reorder_bn ? clip8((s + off) / fac) : clip8(s >> 3));
        }
    }
    for (int n = 0; n < N; n++) {               /* ---- layer 5 */
        int32_t z = p->l5_bias[n];
        for (int k = 0; k < F2; k++)
            for (int q = 0; q < T64; q++)
                z += (int32_t)y4[(size_t)k * T64 + q] * p->l5_weight[((size_t)n * F2 + k) * T64A + q];
        out[n] = (int8_t)clip8(z / p->l5_factor);
    }
    if (taps) {
        for (int k = 0; k < F2; k++) memcpy(taps + (size_t)k * T, y1 + (size_t)k * (T + 63) + 31, T);
        taps += (size_t)F2 * T;
        for (int k = 0; k < F2; k++) memcpy(taps + (size_t)k * T8, y2 + (size_t)k * (T8 + 15) + 7, T8);
        taps += (size_t)F2 * T8;
        memcpy(taps, y3, (size_t)F2 * T8);
        taps += (size_t)F2 * T8;
        memcpy(taps, y4, (size_t)F2 * T64);
    }
}

size_t wolf_port_tap_bytes(int T, int F2) { return (size_t)F2 * (T + 2 * (T / 8) + T / 64); }

/* x [B][C][T] int8 -> out [B][N] int8.  threads <= 0: all cores.  Returns threads used. */
int wolf_port_forward(const wolf_port_net* p, const int8_t* x, long B, int reorder_bn, int8_t* out,
                      int threads, int8_t* taps) {
    const size_t scratch_bytes = (size_t)p->F2 * ((p->T + 63) + (p->T / 8 + 15) + p->T / 8 + p->T / 64) + 64;
    const size_t tapb = wolf_port_tap_bytes(p->T, p->F2);
    int used = 1;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
    used = threads;
#pragma omp parallel num_threads(threads)
#endif
    {
        int8_t* scratch = (int8_t*)malloc(scratch_bytes);
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
        for (long b = 0; b < B; b++)
            forward_one(p, x + (size_t)b * p->C * p->T, reorder_bn, out + (size_t)b * p->N, scratch,
                        taps ? taps + (size_t)b * tapb : NULL);
        free(scratch);
    }
    return used;
}
