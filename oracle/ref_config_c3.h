// Pre-included (nvcc --pre-include) when building the 3-channel variant of the
// reference for the SH config: it pre-empts the include guard of the
// reference's cuda_rasterizer/config.h (which hard-codes NUM_CHANNELS 4,
// config.h:15) without touching or copying any reference source.
// CUB is pulled in first because its histogram headers use NUM_CHANNELS as a
// template parameter name and must be parsed before the macro exists (in the
// reference's own include order config.h also comes after <cub/cub.cuh>,
// rasterizer_impl.cu:20-29).
#ifndef CUDA_RASTERIZER_CONFIG_H_INCLUDED
#ifdef __CUDACC__
#include <cub/cub.cuh>
#include <cub/device/device_radix_sort.cuh>
#endif
#define CUDA_RASTERIZER_CONFIG_H_INCLUDED
#define NUM_CHANNELS 3
#define BLOCK_X 16
#define BLOCK_Y 16
#endif
