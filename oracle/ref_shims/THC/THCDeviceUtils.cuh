// Shim for <THC/THCDeviceUtils.cuh> (deform_psroi_pooling_cuda.cu:20).
#pragma once
#include <ATen/cuda/DeviceUtils.cuh>
