// Host-visible launchers of the CUDA kernels (internal to libmibmi.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mibmi {

struct GenericNet;
struct FusedNet;
struct PipeNet;

size_t generic_smem_bytes(int T, int F2);
cudaError_t launch_generic(const GenericNet* net_dev, int T, int F2, const int8_t* x, long long B, int layout,
                           int8_t* logits, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                           cudaStream_t stream);
cudaError_t launch_quantize(const float* x, long long n, float s, int8_t* q, cudaStream_t stream);
cudaError_t launch_dequantize(const int8_t* q, long long n, float s, float* y, cudaStream_t stream);

// Fused kernel: shapes are compile-time instantiations; returns false if (C, T) has none.
bool fused_has_instance(int C, int T);
// ctas_per_sm <= 0: default.  x must be [B][C][T], 16-byte aligned.
cudaError_t launch_fused(const FusedNet* net_dev, int C, int T, const int8_t* x, long long B, int8_t* logits,
                         int num_sms, int ctas_per_sm, cudaStream_t stream);


// Warp-specialised pipeline kernel (mibmi_pipe.cu): the default fused path.  Optional taps
// (y1 [B][16][T], y2/y3 [B][16][T/8], y4 [B][16][T/64]) select a diagnostic instantiation.
bool pipe_has_instance(int C, int T);
cudaError_t launch_pipe(const PipeNet* net_dev, bool fast1, int C, int T, const int8_t* x, long long B,
                        int8_t* logits, int8_t* t1, int8_t* t2, int8_t* t3, int8_t* t4, int num_sms,
                        cudaStream_t stream);

}  // namespace mibmi
