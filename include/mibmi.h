/* mibmi.h -- C-ABI of the B200-native int8 edgeEEGNet / MI-BMInet inference path.
 *
 * Plain C, no torch types.  This is the boundary a maintainer of the reference binds against; it
 * replaces, for the integer-inference hot path only, the reference's
 *
 *   void net_model_compute(const int8_t* p_data, int8_t* p_output);
 *        edge-eegnet_wolf/src/cl/net/model.h:40, implemented in src/cl/net/model.c:42-106
 *
 * whose network is baked in at link time from the generated arrays of
 *   edge-eegnet_wolf/data/gen_net_header.py:109-224  (net.h / net.c).
 * Here the same arrays are handed over at run time (mibmi_params_t), one handle per network, and a
 * call processes a whole batch of independent trials on the GPU.
 *
 * Integer semantics are the reference's: int8 operands, int32 accumulation,
 * requantisation (acc + offset) / factor with C truncating division (divisor of either sign),
 * clip = __CLIP_R(x, 127) = [-128, 127], '>>' arithmetic.
 *
 * All functions return MIBMI_OK (0) or a negative mibmi_status; nothing prints or aborts.  There is
 * no CPU fallback: without a CUDA device every compute entry point returns MIBMI_ERR_CUDA.
 * Distinct handles may be used concurrently from different threads; one handle may be used from
 * several threads as long as each call is given its own stream: the handle is read-only after
 * creation, scratch memory a call needs is allocated and freed in stream order on that call's own
 * stream.  Exceptions: mibmi_set_option, and the *_host entry points, which own staging buffers and
 * serialise on an internal mutex.
 */
#ifndef MIBMI_H
#define MIBMI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MIBMI_VERSION 100 /* 0.1.0 */

#if defined(__GNUC__)
#define MIBMI_API __attribute__((visibility("default")))
#else
#define MIBMI_API
#endif

typedef enum {
    MIBMI_OK = 0,
    MIBMI_ERR_NULL = -1,        /* null pointer argument */
    MIBMI_ERR_SHAPE = -2,       /* unsupported or inconsistent network dimensions */
    MIBMI_ERR_PARAM = -3,       /* invalid parameter value (zero divisor, bad layout / option ...) */
    MIBMI_ERR_ALIGN = -4,       /* device pointer not sufficiently aligned for the selected layout */
    MIBMI_ERR_CUDA = -5,        /* CUDA runtime error (see mibmi_last_cuda_error) or no device */
    MIBMI_ERR_UNSUPPORTED = -6, /* request valid but not available on the forced kernel path */
    MIBMI_ERR_TOO_LARGE = -7    /* request exceeds the fixed scratch of an allocation-free entry point */
} mibmi_status;

/* Input layouts of one batch of trials (int8, values in [-127, 127] as produced by
 * gen_input_header.py:73; -128 is accepted and handled exactly). */
#define MIBMI_LAYOUT_BCT 0  /* [B][C][T]        -- module order, edgeEEGNet.forward x[B,1,C,T]         */
#define MIBMI_LAYOUT_BTC 1  /* [B][T][C_ALIGN]  -- Wolf order of net_model_compute (gen_input_header.py:75-76);       *
                             * read directly by the fused tcgen05 kernels of every shipped shape: rows of 64 / 16 bytes as   *
                             * TMA tiles, rows of 40 / 24 / 20 / 8 bytes (T = 480) as super-rows of 2 or 4 samples, the      *
                             * 24-byte rows of 22 x 1125 expanded in shared memory -- one launch, no scratch; exact generic  *
                             * kernel for every other shape */

/* Network description: the exported arrays of gen_net_header.py (names without the "net_" prefix).
 * All pointers are HOST pointers; mibmi_create copies what it needs. */
typedef struct {
    int32_t C, T;          /* EEG channels, samples per trial                     (NET_C, NET_T)          */
    int32_t F1, D, F2, N;  /* F2 must equal F1*D and D == 1 (golden_model.py:169); N classes               */
    int32_t Nf;            /* temporal filter length, must be 64 (golden_model.py:216-219)                */
    int32_t reorder_bn;    /* 1: REORDER_BN build (Makefile:34, canonical); 0: plain variant (follows the C); both fused */

    const int32_t* l1_factor;          /* [F2]            gen_net_header.py:109 */
    const int32_t* l1_offset;          /* [F2]            :110                  */
    const int8_t*  l1_weight_align;    /* [F2][C_ALIGN]   :114 (dot-product order, rows padded to 4)     */
    const int32_t* l2_factor;          /* [F2]            :149 (already x8 for the pooling)              */
    const int32_t* l2_offset;          /* [F2]            :150                  */
    const int8_t*  l2_weight_reverse;  /* [F2][64]        :154 (torch / cross-correlation order)         */
    int32_t        l3_factor;          /* NET_L3_FACTOR   :175                  */
    const int8_t*  l3_weight;          /* [F2][16]        :177 (flipped: true-convolution order)         */
    const int32_t* l4_factor;          /* [F2]            :194                  */
    const int32_t* l4_offset;          /* [F2]            :195                  */
    const int8_t*  l4_weight;          /* [F2 out][F2 in] :197                  */
    int32_t        l5_factor;          /* NET_L5_FACTOR   :221                  */
    const int8_t*  l5_bias;            /* [N]             :222                  */
    const int8_t*  l5_weight;          /* [N][F2*T64_ALIGN] :224 (each row of T64 padded to 4)           */
} mibmi_params_t;

typedef struct mibmi_net mibmi_net_t; /* opaque */

/* Kernel paths (mibmi_set_option(MIBMI_OPT_FORCE_PATH, ...), mibmi_last_path). */
#define MIBMI_PATH_AUTO 0     /* fused tensor-core kernel when the network qualifies, else generic    */
#define MIBMI_PATH_GENERIC 1  /* exact reference-order kernel: any shape, both reorder_bn modes, taps */
#define MIBMI_PATH_FUSED 2    /* fused sm_100a kernel: tcgen05 / TMEM kernel where the shape has one, else  *
                               * the mma.sync pipeline kernel (magic-number requantisation in both)          */
#define MIBMI_PATH_FUSED_PIPE 5 /* force the warp-specialised mma.sync pipeline kernel (A/B measurements)   */
#define MIBMI_PATH_STREAM 4   /* sliding-window streaming kernel (mibmi_forward_stream_int8)             */

#define MIBMI_STREAM_AUTO 0    /* fused streaming kernel when it applies (hop 16, L % 16 == 0), else gather */
#define MIBMI_STREAM_FUSED 1   /* fused streaming kernel or MIBMI_ERR_UNSUPPORTED                          */
#define MIBMI_STREAM_GATHER 2  /* materialise the windows, then the per-trial path (any hop)               */

#define MIBMI_OPT_FORCE_PATH 1
#define MIBMI_OPT_HOST_CHUNK 3    /* trials per H2D/compute/D2H chunk of the *_host entry point; 0 = default */
#define MIBMI_OPT_STREAM_MODE 4   /* MIBMI_STREAM_* */

MIBMI_API int mibmi_version(void);
MIBMI_API const char* mibmi_strerror(int status);
/* cudaError_t of the most recent failing CUDA call on this handle (0 if none). */
MIBMI_API int mibmi_last_cuda_error(const mibmi_net_t* net);

/* Validate and upload a network to `device` (CUDA ordinal). */
MIBMI_API int mibmi_create(const mibmi_params_t* params, int device, mibmi_net_t** out);
MIBMI_API int mibmi_destroy(mibmi_net_t* net);

MIBMI_API int mibmi_set_option(mibmi_net_t* net, int option, int64_t value);
/* Which path the most recent forward on this handle ran (MIBMI_PATH_GENERIC / MIBMI_PATH_FUSED). */
MIBMI_API int mibmi_last_path(const mibmi_net_t* net);
/* Name of the __global__ function the most recent forward launched, recorded at launch time ("tc_forward_kernel",
 * "tcr_forward_kernel", "pipe_forward_kernel", "stream_forward_kernel", "generic_forward_kernel"; "" before the first
 * call).  Static string. */
MIBMI_API const char* mibmi_last_kernel(const mibmi_net_t* net);
/* 1 if the fused path can run this network (shape supported, requant parameters within the proven
 * range of the magic-number division), else 0 with the reason in *why (static string, may be NULL). */
MIBMI_API int mibmi_fused_supported(const mibmi_net_t* net, const char** why);
/* Kernel launches issued on behalf of this handle since creation (evidence counter). */
MIBMI_API int64_t mibmi_launch_count(const mibmi_net_t* net);

/* Batched forward, device buffers, asynchronous on `stream` (a cudaStream_t, may be NULL).
 *   x_dev      int8, B trials in `layout`.  The fused kernels need it 16-byte aligned; no byte outside
 *              [x_dev, x_dev + B * trial_bytes) is read.  With MIBMI_PATH_AUTO a pointer that is not 16-byte
 *              aligned runs on the exact generic kernel (no alignment requirement); with a fused path forced it
 *              is MIBMI_ERR_ALIGN.
 *   logits_dev int8 [B][N], caller-allocated; 4-byte aligned for the fused kernels
 * Any B >= 0; batches above 2^20 trials are split into several launches.
 * Replaces B calls of net_model_compute (model.c:42). */
MIBMI_API int mibmi_forward_int8(mibmi_net_t* net, const int8_t* x_dev, int64_t B, int layout,
                       int8_t* logits_dev, void* stream);

/* Forward fused with the all-gather of a batch that is sharded over several GPUs: the logits of
 * the B local trials are written to rows [row0, row0 + B) of EVERY buffer out_dev[0 .. n_out-1] (int8 [rows][N];
 * out_dev is a HOST array of device pointers: this rank's buffer and the peer-mapped buffers of the other ranks,
 * e.g. CUDA IPC / symmetric memory over NVLink).  The fused kernel stores the 4 bytes of a trial to every
 * destination from its epilogue -- no collective is launched; other paths copy peer to peer on `stream`.
 * The caller synchronises the ranks (one barrier) before anyone reads a gathered buffer.  n_out <= 16. */
MIBMI_API int mibmi_forward_int8_gather(mibmi_net_t* net, const int8_t* x_dev, int64_t B, int layout,
                                        int8_t* const* out_dev, int32_t n_out, int64_t row0, void* stream);

/* Same, plus the inter-layer activations (the taps of QuantLab/export_net_data.py:112-167 in integer
 * form).  Any of the tap pointers may be NULL.  Runs the generic path unless MIBMI_PATH_FUSED is forced
 * (the fused kernel then stores the taps from its registers: a diagnostic instantiation).
 *   y1 [B][F2][T], y2 [B][F2][T/8], y3 [B][F2][T/8], y4 [B][F2][T/64]  (int8, device) */
MIBMI_API int mibmi_forward_int8_taps(mibmi_net_t* net, const int8_t* x_dev, int64_t B, int layout,
                            int8_t* logits_dev, int8_t* y1, int8_t* y2, int8_t* y3, int8_t* y4,
                            void* stream);

/* End-to-end convenience for callers that hold HOST buffers (what a cgo / ctypes caller of the
 * reference has): chunked H2D copy -> forward -> D2H copy, double-buffered on two internal streams
 * through pinned staging memory.  Synchronous: returns when logits_host is complete. */
MIBMI_API int mibmi_forward_int8_host(mibmi_net_t* net, const int8_t* x_host, int64_t B, int layout,
                            int8_t* logits_host);

/* Sliding-window scoring of continuous recordings (BASELINE config 5).  x_dev int8 [S][C][L]: S independent
 * streams (subjects) of L samples, 16-byte aligned; window w of stream s is x[s][:, w*hop : w*hop + T] and is
 * scored exactly as one call of net_model_compute (model.c:42) on it -- per-window zero padding included.
 *   logits_dev int8 [S][W][N],  W = mibmi_stream_windows(net, L, hop) = (L - T) / hop + 1
 * With hop == 16 and L % 16 == 0 the fused streaming kernel computes layer 1 once per sample and the interior
 * of layer 2 once per 8 samples instead of once per window; any other hop materialises the windows in a
 * stream-ordered scratch buffer and runs the per-trial path.  Asynchronous on `stream`. */
MIBMI_API int64_t mibmi_stream_windows(const mibmi_net_t* net, int64_t L, int32_t hop);
MIBMI_API int mibmi_stream_supported(const mibmi_net_t* net, const char** why);
MIBMI_API int mibmi_forward_stream_int8(mibmi_net_t* net, const int8_t* x_dev, int64_t S, int64_t L, int32_t hop,
                                        int8_t* logits_dev, void* stream);
/* Same on the fused streaming kernel, plus its intermediate results (any pointer may be NULL):
 *   y2_stream [S][16][2 (W-1) + T/8]  pooled layer-2 samples of the stream (position j is written where it is
 *                                     interior to some window: y2 of window w, sample p in [4, T/8-5], is [2 w + p])
 *   y2_edges  [S][W][16][8]           y2 samples 0..3 and T/8-4 .. T/8-1 of every window
 *   y3 [S][W][16][T/8], y4 [S][W][16][T/64] */
MIBMI_API int mibmi_forward_stream_int8_taps(mibmi_net_t* net, const int8_t* x_dev, int64_t S, int64_t L, int32_t hop,
                                             int8_t* logits_dev, int8_t* y2_stream, int8_t* y2_edges, int8_t* y3,
                                             int8_t* y4, void* stream);

/* Allocation-free variant of the any-hop (window-materialising) form of mibmi_forward_stream_int8.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  The windows are materialised in a file-scope `static __device__` array of
 * MIBMI_NOALLOC_SCRATCH_BYTES (32 MiB) that every GPU which loads the library holds until exit; the call itself performs no
 * cudaMalloc / cudaFree / stream-ordered allocation.  Same arguments, results and alignment rules as mibmi_forward_stream_int8
 * (any hop >= 1; always the gather path: one gather_windows launch + one per-trial forward launch).
 *   S * W * C * T > MIBMI_NOALLOC_SCRATCH_BYTES  ->  MIBMI_ERR_TOO_LARGE (nothing is launched; split the request)
 * All calls on a device share the one array, which keeps the previous call's contents (every byte the forward reads is
 * rewritten by the gather of the same call, nothing has to start at zero): calls that overlap on the GPU (different streams /
 * threads / handles on the same device) RACE -- serialise them.  Asynchronous on `stream`. */
#define MIBMI_NOALLOC_SCRATCH_BYTES (32 * 1024 * 1024)
MIBMI_API int mibmi_forward_stream_int8_noalloc(mibmi_net_t* net, const int8_t* x_dev, int64_t S, int64_t L, int32_t hop,
                                                int8_t* logits_dev, void* stream);

/* Input quantisation pre-op (gen_input_header.py:73, STEActivation with floorToZero,
 * QuantLab/quantlab/indiv/ste_ops.py:88-102):  q = trunc(clip(x / abs_max, -1, 1) * 127).
 * x_dev float32 [n] -> q_dev int8 [n] (same element order). */
MIBMI_API int mibmi_quantize_input_f32(const float* x_dev, int64_t n, float abs_max, int8_t* q_dev, void* stream);
/* Output de-quantisation post-op (functional.py:338-353): y = q / 127 * abs_max. */
MIBMI_API int mibmi_dequantize_logits_f32(const int8_t* q_dev, int64_t n, float abs_max, float* y_dev, void* stream);

/* Diagnostics: the (magic, shift) pair the fused path uses for a positive divisor d >= 4, such that
 * trunc(x / d) == (mulhi32(x, magic) >> shift) + (x < 0) for |x| < 2^29.  MIBMI_ERR_PARAM if d < 4. */
MIBMI_API int mibmi_debug_magic(int32_t d, int32_t* magic, int32_t* shift);
/* Diagnostics: the magic of the shift-free layer-1 form clip8(mulhi32(x, magic) + (x < 0)) == clip8(trunc(x / d)) for every
 * |x| <= xmax (the bound mibmi_create proves per filter: 128 * sum|w| + |offset|).  MIBMI_ERR_UNSUPPORTED when the form is
 * not exact for this (d, xmax): the layer then uses the shifted form of mibmi_debug_magic. */
MIBMI_API int mibmi_debug_magic32(int32_t d, int64_t xmax, int32_t* magic);
/* Diagnostics (no GPU needed): writes the packed network the fused kernels (MIBMI_PATH_FUSED) read into out[cap]
 * and returns its size; out == NULL only queries the size.  MIBMI_ERR_UNSUPPORTED (+ *why) if the fused
 * path cannot run this network. */
MIBMI_API int64_t mibmi_debug_pack_pipe(const mibmi_params_t* params, void* out, int64_t cap, const char** why);

/* ---------------------------------------------------------------------------------------------------------
 * EEGNet / Q-EEGNet layer order: temporal conv (1, Nf) over every EEG channel first, then the
 * depthwise spatial conv -- replaces QuantLab/quantlab/PhysionetMMMI/EEGNet/eegnet.py:137-168 (EEGNet.forward) for
 * integer inference.  The reference tree has no integer model of this topology; semantics = the edgeEEGNet integer
 * pipeline's (golden_model.py:153-379, REORDER_BN form) applied to this layer order, see oracle/eegnet_np.py.
 * All weights in torch (cross-correlation) order, HOST pointers, copied by create.
 * Kernels (same bytes from all of them; mibmi_eegnet_last_kernel names the one that ran): F1 = 8 filters of Nf = 128 taps with
 * C <= 64 (the QuantLab EEGNet configuration) run both contractions on tcgen05.mma with TMEM accumulators; Nf a multiple of 32
 * otherwise on mma.sync; anything else on dp4a. */
typedef struct {
    int32_t C, T;              /* EEG channels, samples per trial                                   */
    int32_t F1, D, F2, N;      /* F1 % 4 == 0, F1 <= 16, D in {1, 2}, F2 <= 32, N <= 8             */
    int32_t Nf;                /* temporal filter length (conv1 kernel (1, Nf)), multiple of 4, <= 128; pad Nf/2-1 | Nf/2 */
    const int8_t*  l1_weight;  /* [F1][Nf]        conv1 (eegnet.py:101-102)                          */
    const int32_t* l1_factor;  /* [F1]            batch_norm1 folded, quant2 (pool = 1)              */
    const int32_t* l1_offset;  /* [F1]                                                              */
    const int8_t*  l2_weight;  /* [F1*D][C]       conv2, depthwise spatial, groups = F1 (eegnet.py:105) */
    const int32_t* l2_factor;  /* [F1*D]          batch_norm2 + ReLU + pool 8 folded (x8), quant3    */
    const int32_t* l2_offset;  /* [F1*D]                                                            */
    const int8_t*  l3_weight;  /* [F1*D][16]      sep_conv1 (eegnet.py:113)                          */
    int32_t        l3_factor;
    const int8_t*  l4_weight;  /* [F2][F1*D]      sep_conv2 (eegnet.py:115)                          */
    const int32_t* l4_factor;  /* [F2]            batch_norm3 + ReLU + pool 8 folded, quant5         */
    const int32_t* l4_offset;  /* [F2]                                                              */
    const int8_t*  l5_weight;  /* [N][F2*T64]     fc, flatten order (k, q)                           */
    const int8_t*  l5_bias;    /* [N]                                                               */
    int32_t        l5_factor;
} mibmi_eegnet_params_t;

typedef struct mibmi_eegnet mibmi_eegnet_t; /* opaque */

MIBMI_API int mibmi_eegnet_create(const mibmi_eegnet_params_t* params, int device, mibmi_eegnet_t** out);
MIBMI_API int mibmi_eegnet_destroy(mibmi_eegnet_t* net);
/* x_dev: int8 [B][C][T]; logits_dev: int8 [B][N]; asynchronous on `stream`. */
MIBMI_API int mibmi_eegnet_forward_int8(mibmi_eegnet_t* net, const int8_t* x_dev, int64_t B, int8_t* logits_dev, void* stream);
/* Same with per-layer taps (any may be NULL): y2 [B][F1*D][T/8], y3 [B][F1*D][T/8], y4 [B][F2][T/64]. */
MIBMI_API int mibmi_eegnet_forward_int8_taps(mibmi_eegnet_t* net, const int8_t* x_dev, int64_t B, int8_t* logits_dev,
                                             int8_t* y2, int8_t* y3, int8_t* y4, void* stream);
MIBMI_API int64_t mibmi_eegnet_launch_count(const mibmi_eegnet_t* net);
/* Name of the __global__ function the most recent forward of this handle launched ("eegnet_tc_kernel": both contractions on
 * tcgen05.mma with TMEM accumulators, F1 = 8 and Nf = 128; "eegnet_mma_kernel": mma.sync; "eegnet_forward_kernel": dp4a). */
MIBMI_API const char* mibmi_eegnet_last_kernel(const mibmi_eegnet_t* net);

#ifdef __cplusplus
}
#endif
#endif /* MIBMI_H */
