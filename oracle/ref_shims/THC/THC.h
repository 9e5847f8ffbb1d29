// Shim for the header the reference's deform_psroi_pooling_cuda.cu:18 includes (THC was removed from PyTorch).
// Test infrastructure: lets oracle/refenv.py compile the UNMODIFIED reference sources under /root/reference/deform/src.
#pragma once
#include <c10/cuda/CUDAException.h>
#define THCudaCheck(expr) C10_CUDA_CHECK(expr)
template <typename T> inline T THCCeilDiv(T a, T b) { return (a + b - 1) / b; }
