// Host shim so that the reference's gpu-version/*.cuh headers compile as plain C++
// (test infrastructure only; see oracle/README.md).  Placed FIRST on the include path
// when building oracle/_ref/ref_gpu_harness.  No reference code is copied here.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>

#define __device__
#define __host__
#define __global__

typedef int cudaError_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
enum cudaLimit { cudaLimitStackSize };

static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = malloc(n ? n : 1); return 0; }
static inline cudaError_t cudaMallocManaged(void **p, size_t n) { *p = malloc(n ? n : 1); return 0; }
template <class T> static inline cudaError_t cudaMalloc(T **p, size_t n) { return cudaMalloc((void **)p, n); }
template <class T> static inline cudaError_t cudaMallocManaged(T **p, size_t n) { return cudaMallocManaged((void **)p, n); }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaFree(void *p) { free(p); return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaDeviceReset() { return 0; }
static inline cudaError_t cudaPeekAtLastError() { return 0; }
static inline cudaError_t cudaDeviceSetLimit(cudaLimit, size_t) { return 0; }
static inline const char *cudaGetErrorString(cudaError_t) { return ""; }
