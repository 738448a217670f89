// rt_noalloc.cu — rt_tonemap_noalloc: the tonemap + read-back of rt_download for callers that must not allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  rt_download tonemaps into a staging buffer it
// allocates with cudaMalloc on first use (rt_ctx::d_rgb8); here the staging buffer is a file-scope __device__ array of
// RT_NOALLOC_MAX_BYTES that every GPU loading the library holds until the process exits.  The entry point calls no cudaMalloc /
// cudaFree / allocating library routine: one kernel launch into the static array, one cudaMemcpyAsync from it to the caller's
// host buffer, one stream synchronisation.
//
// The array's NAME is a host-side stand-in in host code, not a device address: its device address is looked up once per device
// (with that device current) through cudaGetSymbolAddress, and that pointer is what the kernel and the copy receive.
//
// Sharing: every call on a device uses the same array and it keeps the previous call's bytes.  Nothing needs to start at zero —
// the kernel overwrites bytes [0, width*height*3) and only those are copied out — but two calls that overlap on the same GPU
// (different streams, different host threads) race on it.  Serialise them.
#include "rt_ctx.h"
#include <mutex>

static __device__ __align__(256) unsigned char g_noalloc_scratch[RT_NOALLOC_MAX_BYTES];

// write_image (PNG: sum/spp, clamp, x256, no gamma — color.cuh:15-35) / write_color (PPM: sqrt gamma — color.cuh:70-95), rows
// flipped so that row 0 of the output is the top row: the same arithmetic as k_tonemap (rt_kernels.cu)
__global__ void __launch_bounds__(256) k_tonemap_noalloc(const float *__restrict__ accum, unsigned char *__restrict__ rgb8, int W, int H,
                                                         int spp, int gamma) {
    const int n = W * H * 3;
    const float fspp = (float)spp, scale = 1.0f / fspp;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int c = i % 3, px = i / 3, row = px / W, col = px - row * W;
        float x = accum[3 * ((H - 1 - row) * W + col) + c];
        x = gamma ? sqrtf(x * scale) : x / fspp;
        x = x < 0.0f ? 0.0f : (x > 0.999f ? 0.999f : x);
        rgb8[i] = (unsigned char)(int)(256.0f * x);
    }
}

namespace {
struct DevSlot { std::once_flag once; cudaError_t err = cudaSuccess; unsigned char *ptr = nullptr; int sms = 0; };
DevSlot g_slot[RT_MAX_DEVICES];
}

extern "C" int rt_tonemap_noalloc(int device, void *stream, const float *accum_dev, int width, int height, int total_spp, int gamma,
                                  unsigned char *rgb8_host) {
    if (!accum_dev || !rgb8_host) return rt_fail(nullptr, RT_ERR_INVALID, "rt_tonemap_noalloc: NULL buffer");
    if (width < 1 || height < 1 || total_spp < 1) return rt_fail(nullptr, RT_ERR_INVALID, "rt_tonemap_noalloc: width, height and total_spp must be >= 1");
    const size_t bytes = (size_t)width * (size_t)height * 3;
    if (bytes > (size_t)RT_NOALLOC_MAX_BYTES)
        return rt_fail(nullptr, RT_ERR_INVALID, "rt_tonemap_noalloc: image of " + std::to_string(bytes) + " bytes exceeds the static scratch of " +
                                                    std::to_string((size_t)RT_NOALLOC_MAX_BYTES) + " bytes (RT_NOALLOC_MAX_BYTES); use rt_download");
    if (device < 0 || device >= RT_MAX_DEVICES) return rt_fail(nullptr, RT_ERR_INVALID, "rt_tonemap_noalloc: bad device ordinal");
    int prev = -1;
    cudaGetDevice(&prev);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return rt_fail(nullptr, RT_ERR_CUDA, std::string("rt_tonemap_noalloc: cudaSetDevice: ") + cudaGetErrorString(e));
    DevSlot &slot = g_slot[device];
    std::call_once(slot.once, [&] {          // once per device, with that device current
        slot.err = cudaGetSymbolAddress((void **)&slot.ptr, g_noalloc_scratch);
        if (slot.err == cudaSuccess) slot.err = cudaDeviceGetAttribute(&slot.sms, cudaDevAttrMultiProcessorCount, device);
    });
    int rc = RT_OK;
    if (slot.err != cudaSuccess || !slot.ptr) {
        rc = rt_fail(nullptr, RT_ERR_CUDA, std::string("rt_tonemap_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(slot.err));
    } else {
        cudaStream_t s = (cudaStream_t)stream;
        k_tonemap_noalloc<<<slot.sms * 8, 256, 0, s>>>(accum_dev, slot.ptr, width, height, total_spp, gamma);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpyAsync(rgb8_host, slot.ptr, bytes, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) rc = rt_fail(nullptr, RT_ERR_CUDA, std::string("rt_tonemap_noalloc: ") + cudaGetErrorString(e));
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    return rc;
}
