// Build-time stand-in for the reference's CuMatrix/Buffers/ManagedBuffer.h (device-buffer RAII wrappers that need a
// CUDA device and do not compile under GCC). The reference's __host__ __device__ math only needs the basic typedefs.
#pragma once
#include <cstdint>
#include <cmath>
#include <memory>
#include "cuda_runtime.h"
#include "CuMatrix/MatrixOps/CuMatrixDefs.h"
