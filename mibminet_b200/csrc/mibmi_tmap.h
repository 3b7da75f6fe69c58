// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace mibmi {

typedef CUresult (*tmap_encode_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
tmap_encode_t tmap_encoder();   // nullptr if the driver does not provide it (defined in mibmi_pipe.cu)

}  // namespace mibmi
