// Host-visible launchers of the CUDA kernels (internal to libmibmi.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mibmi {

struct GenericNet;
struct PipeNet;
struct StreamNet;

size_t generic_smem_bytes(int T, int F2);
cudaError_t launch_generic(const GenericNet* net_dev, int T, int F2, const int8_t* x, long long B, int layout,
                           int8_t* logits, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                           cudaStream_t stream);
cudaError_t launch_quantize(const float* x, long long n, float s, int8_t* q, cudaStream_t stream);
cudaError_t launch_dequantize(const int8_t* q, long long n, float s, float* y, cudaStream_t stream);

// Warp-specialised pipeline kernel (mibmi_pipe.cu): the default fused path.  Optional taps
// (y1 [B][16][T], y2/y3 [B][16][T/8], y4 [B][16][T/64]) select a diagnostic instantiation.
// Extra destinations of the logits (peer-mapped buffers of the other ranks): the all-gather of a sharded batch
// is done by the kernel's own stores over NVLink.
constexpr int kMaxPeers = 15;
struct PeerOut {
    int8_t* p[kMaxPeers];
    int n;
};
bool pipe_has_instance(int C, int T);
cudaError_t launch_pipe(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B,
                        int8_t* logits, const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                        cudaStream_t stream);

// Tensor-core-resident kernel (mibmi_tc.cu): layers 1 and 2 on tcgen05.mma with TMEM accumulators; the default
// fused path for shapes with 16-byte-aligned rows.  Same PipeNet, same optional taps.
bool tc_has_instance(int C, int T);
cudaError_t launch_tc(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B, int8_t* logits,
                      const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream,
                      bool plain = false);      // plain: the REORDER_BN-off variant (PipeNet::plain), module order only

// The same kernel fed from the Wolf order [B][T][C_ALIGN] of net_model_compute (K-major layer-1 operand through a 3-D tensor map).
bool tc_has_wolf_instance(int C, int T);
cudaError_t launch_tc_wolf(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B, int8_t* logits,
                           const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream);

// Same formulation for trials whose rows are not 16-byte aligned (mibmi_tcr.cu: realigning warps instead of TMA tiles,
// MMA tiles that straddle trials): 22 x 1125.
bool tcr_has_instance(int C, int T);
// layout 0: x [B][C][T] (rows realigned by funnel shifts); layout 1: Wolf order x [B][T][C_ALIGN] (rows of 24 bytes expanded
// to the K-major 32-byte rows of the layer-1 operand inside the kernel -- no scratch, no second pass over HBM)
bool tcr_has_wolf_instance(int C, int T);
cudaError_t launch_tcr(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B, int layout, int8_t* logits,
                       const PeerOut& peers, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream,
                       bool plain = false);

// Wolf order [n][T][C_ALIGN] -> module order [n][C][T] (mibmi_reorder.cu) for rows of 8 / 20 / 24 / 40 bytes; y 4-byte aligned.
bool reorder_has_instance(int C, int T);
cudaError_t launch_wolf_reorder(const int8_t* x, int8_t* y, long long n, int C, int T, int num_sms, cudaStream_t stream);

// Sliding-window streaming kernel (mibmi_stream.cu): x [S][C][L] -> logits [S][W][4], W = (L - T) / 16 + 1.
// Optional taps: t2i [S][16][2 (W-1) + T/8] stream-level pooled layer-2 samples (interior positions only),
// t2e [S][W][16][8] the 4 + 4 edge samples of every window, t3 [S][W][16][T/8], t4 [S][W][16][T/64].
bool stream_has_instance(int C, int T);
cudaError_t launch_stream(const StreamNet* net_dev, int fast /* bit 0: layer 1, bit 1: layer 3 in the no-shift form */, int C, int T, const int8_t* x, long long S, long long L,
                          int8_t* logits, int8_t* t2i, int8_t* t2e, int8_t* t3, int8_t* t4, int num_sms, cudaStream_t stream);
// Fallback: windows first .. first + nw - 1 (global index subject * W + w) of the streams -> out [nw][C][T]
cudaError_t launch_gather_windows(const int8_t* x, long long L, int C, int T, int hop, long long W, long long first,
                                  long long nw, int8_t* out, cudaStream_t stream);

}  // namespace mibmi
