// Shim for <THC/THCAtomics.cuh> (deform/src/cuda/*_im2col_cuda.cuh:9): atomicAdd overloads now live in ATen.
#pragma once
#include <ATen/cuda/Atomic.cuh>
