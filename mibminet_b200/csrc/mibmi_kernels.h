// Host-visible launchers of the CUDA kernels (internal to libmibmi.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mibmi {

struct GenericNet;
struct FusedNet;

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

}  // namespace mibmi
