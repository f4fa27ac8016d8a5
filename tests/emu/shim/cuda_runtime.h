// tests/emu/shim/cuda_runtime.h — TEST INFRASTRUCTURE ONLY.
//
// A stand-in for <cuda_runtime.h> that lets g++ compile the product's CUDA sources (csrc/*.cu, *.cuh) for the
// HOST, so that the kernels' logic can be checked against the oracle in a container without a GPU
// (tests/test_emu_cpu.py).  It is an SIMT emulator, not a port: every CUDA thread of a block is a fiber
// (ucontext), the 32 lanes of a warp meet at each *_sync intrinsic (ballot, shuffle, match, syncwarp) exactly as
// converged lanes do on the device, __shared__ variables are block-wide statics, blocks run one after another.
// Device memory is host memory, streams and events are no-ops (everything is synchronous).
//
// Nothing under myraytracer_b200/ links or loads this: librtcuda_emu.so is built by tests/emu/Makefile into
// tests/emu/_build/ and loaded by the emulation tests alone.  The product library has no CPU path.
#pragma once
#ifndef RT_EMU
#error "the emulation shim is for -DRT_EMU test builds only"
#endif

#include <fenv.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <chrono>
#include <cmath>
#include <functional>

using std::isfinite;

// ---- keywords ---------------------------------------------------------------------------------------------------
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline __attribute__((always_inline))
#define __launch_bounds__(...)
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))

// ---- vector types -----------------------------------------------------------------------------------------------
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) int4 { int x, y, z, w; };
struct alignas(8) int2 { int x, y; };
struct alignas(8) uint2 { unsigned x, y; };
struct uchar4 { unsigned char x, y, z, w; };
struct alignas(16) ulonglong2 { unsigned long long x, y; };
struct uint3 { unsigned x, y, z; };
struct dim3 { unsigned x, y, z; dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {} };
static inline int2 make_int2(int x, int y) { int2 r; r.x = x; r.y = y; return r; }
static inline uint2 make_uint2(unsigned x, unsigned y) { uint2 r; r.x = x; r.y = y; return r; }
static inline ulonglong2 make_ulonglong2(unsigned long long x, unsigned long long y) { ulonglong2 r; r.x = x; r.y = y; return r; }

// ---- the emulator (tests/emu/emu_runtime.cpp) ---------------------------------------------------------------------
namespace rtemu {
extern uint3 g_threadIdx, g_blockIdx;
extern dim3 g_blockDim, g_gridDim;
extern uintptr_t g_sharedBase;                  // origin of the emulated 32-bit shared window (an address in .bss)
void launch(dim3 grid, dim3 block, const std::function<void()>& body);
void warpBarrier();
void blockBarrier();
uint64_t* exchangeSlot(int lane);               // this warp's slot for `lane` in the current exchange buffer
int warpLanes();                                // lanes that exist in the calling thread's warp
void nextExchange();                            // flip this lane's exchange buffer (once per collective)
extern unsigned long long g_kernelLaunches;
}  // namespace rtemu
#define threadIdx (rtemu::g_threadIdx)
#define blockIdx (rtemu::g_blockIdx)
#define blockDim (rtemu::g_blockDim)
#define gridDim (rtemu::g_gridDim)

// ---- warp collectives: all lanes of the warp must call them (converged full-mask use, as in the kernels) ---------
namespace rtemu {
template <class T>
inline T collectiveRead(T mine, int src) {
    static_assert(sizeof(T) <= 8, "exchange slot is 8 bytes");
    const int lane = (int)(g_threadIdx.x & 31u);
    nextExchange();
    uint64_t raw = 0;
    memcpy(&raw, &mine, sizeof(T));
    *exchangeSlot(lane) = raw;
    warpBarrier();
    T out = mine;
    if (src >= 0 && src < warpLanes()) memcpy(&out, exchangeSlot(src), sizeof(T));
    return out;
}
}  // namespace rtemu
template <class T> inline T __shfl_sync(unsigned, T v, int src) { return rtemu::collectiveRead(v, src & 31); }
template <class T> inline T __shfl_down_sync(unsigned, T v, unsigned off) {
    const int src = (int)(threadIdx.x & 31u) + (int)off;
    return rtemu::collectiveRead(v, src < 32 ? src : -1);
}
template <class T> inline T __shfl_up_sync(unsigned, T v, unsigned off) {
    const int src = (int)(threadIdx.x & 31u) - (int)off;
    return rtemu::collectiveRead(v, src);
}
inline unsigned __ballot_sync(unsigned, int pred) {
    const int lane = (int)(threadIdx.x & 31u);
    rtemu::nextExchange();
    *rtemu::exchangeSlot(lane) = pred ? 1u : 0u;
    rtemu::warpBarrier();
    unsigned m = 0;
    for (int l = 0; l < rtemu::warpLanes(); l++) if (*rtemu::exchangeSlot(l)) m |= 1u << l;
    return m;
}
inline unsigned __match_any_sync(unsigned, int v) {
    const int lane = (int)(threadIdx.x & 31u);
    rtemu::nextExchange();
    *rtemu::exchangeSlot(lane) = (uint64_t)(uint32_t)v;
    rtemu::warpBarrier();
    unsigned m = 0;
    for (int l = 0; l < rtemu::warpLanes(); l++) if (*rtemu::exchangeSlot(l) == (uint64_t)(uint32_t)v) m |= 1u << l;
    return m;
}
inline int __reduce_min_sync(unsigned, int v) {
    const int lane = (int)(threadIdx.x & 31u);
    rtemu::nextExchange();
    *rtemu::exchangeSlot(lane) = (uint64_t)(uint32_t)v;
    rtemu::warpBarrier();
    int m = v;
    for (int l = 0; l < rtemu::warpLanes(); l++) { const int o = (int)(uint32_t)*rtemu::exchangeSlot(l); if (o < m) m = o; }
    return m;
}
inline int __reduce_max_sync(unsigned, int v) {
    const int lane = (int)(threadIdx.x & 31u);
    rtemu::nextExchange();
    *rtemu::exchangeSlot(lane) = (uint64_t)(uint32_t)v;
    rtemu::warpBarrier();
    int m = v;
    for (int l = 0; l < rtemu::warpLanes(); l++) { const int o = (int)(uint32_t)*rtemu::exchangeSlot(l); if (o > m) m = o; }
    return m;
}
inline void __syncwarp(unsigned = 0xffffffffu) { rtemu::warpBarrier(); }
inline void __syncthreads() { rtemu::blockBarrier(); }

// ---- scalar intrinsics --------------------------------------------------------------------------------------------
template <class T> inline T __ldg(const T* p) { return *p; }
inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
// one rounding, towards -inf / +inf (FFMA.RM / FFMA.RP): fmaf under the directed rounding mode
inline float __fmaf_rd(float a, float b, float c) { const int o = fegetround(); fesetround(FE_DOWNWARD); volatile float r = fmaf(a, b, c); fesetround(o); return r; }
inline float __fmaf_ru(float a, float b, float c) { const int o = fegetround(); fesetround(FE_UPWARD); volatile float r = fmaf(a, b, c); fesetround(o); return r; }
inline float __fdividef(float a, float b) { return a / b; }
inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
inline float __double2float_rd(double x) { float f = (float)x; return (double)f > x ? nextafterf(f, -INFINITY) : f; }
inline float __double2float_ru(double x) { float f = (float)x; return (double)f < x ? nextafterf(f, INFINITY) : f; }
inline double __longlong_as_double(long long v) { double d; memcpy(&d, &v, 8); return d; }
inline long long __double_as_longlong(double d) { long long v; memcpy(&v, &d, 8); return v; }
inline float __uint_as_float(unsigned u) { float f; memcpy(&f, &u, 4); return f; }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline int __clz(int v) { return v == 0 ? 32 : __builtin_clz((unsigned)v); }
inline size_t __cvta_generic_to_shared(const void* p) { return (size_t)((uintptr_t)p - rtemu::g_sharedBase); }
// blocks run one after another and fibers switch only at collectives: plain read-modify-write is atomic here
inline int atomicAdd(int* p, int v) { int o = *p; *p = o + v; return o; }
inline unsigned atomicAdd(unsigned* p, unsigned v) { unsigned o = *p; *p = o + v; return o; }
inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; *p = o + v; return o; }
inline int atomicMax(int* p, int v) { int o = *p; if (v > o) *p = v; return o; }
inline unsigned long long atomicMax(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; if (v > o) *p = v; return o; }
inline unsigned long long atomicMin(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; if (v < o) *p = v; return o; }

// ---- runtime API: device memory is host memory, everything is synchronous ----------------------------------------
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorInvalidValue = 1, cudaErrorMemoryAllocation = 2, cudaErrorNotReady = 600, cudaErrorNotSupported = 801,
       cudaErrorUnknown = 999 };
typedef struct rtemuStream* cudaStream_t;
struct rtemuEvent { double ms; };
typedef rtemuEvent* cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum { cudaEventDisableTiming = 2, cudaStreamNonBlocking = 1, cudaIpcMemLazyEnablePeerAccess = 1, cudaHostAllocPortable = 1 };
struct cudaDeviceProp { int multiProcessorCount; char name[64]; };
struct cudaIpcMemHandle_t { char reserved[64]; };

inline const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "emulated CUDA error"; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
// RTEMU_DEVICES=n: n emulated devices (all of them this host's memory) for the multi-device paths of the C ABI
inline cudaError_t cudaGetDeviceCount(int* n) { const char* e = getenv("RTEMU_DEVICES"); *n = e ? atoi(e) : 1; if (*n < 1) *n = 1; return cudaSuccess; }
inline cudaError_t cudaDeviceCanAccessPeer(int* can, int, int) { *can = 1; return cudaSuccess; }
inline cudaError_t cudaDeviceEnablePeerAccess(int, unsigned) { return cudaSuccess; }
inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int) { memset(p, 0, sizeof *p); p->multiProcessorCount = 1; return cudaSuccess; }
inline cudaError_t cudaMalloc(void** p, size_t n) { *p = calloc(n ? n : 1, 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
inline cudaError_t cudaFree(void* p) { free(p); return cudaSuccess; }
template <class T> inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) { *p = (void*)&symbol; return cudaSuccess; }   // device memory is host memory
inline cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
inline cudaError_t cudaFreeHost(void* p) { free(p); return cudaSuccess; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyPeerAsync(void* d, int, const void* s, int, size_t n, cudaStream_t = nullptr) { memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpy2DAsync(void* d, size_t dpitch, const void* s, size_t spitch, size_t width, size_t height, cudaMemcpyKind,
                                     cudaStream_t = nullptr) {
    for (size_t r = 0; r < height; r++) memmove((char*)d + r * dpitch, (const char*)s + r * spitch, width);
    return cudaSuccess;
}
inline cudaError_t cudaHostAlloc(void** p, size_t n, unsigned) { *p = calloc(n ? n : 1, 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
inline cudaError_t cudaMemset(void* d, int v, size_t n) { memset(d, v, n); return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { memset(d, v, n); return cudaSuccess; }
inline double rtemuNowMs() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new rtemuEvent{0.0}; return cudaSuccess; }
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { return cudaEventCreate(e); }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = nullptr) { e->ms = rtemuNowMs(); return cudaSuccess; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventQuery(cudaEvent_t) { return cudaSuccess; }          // synchronous emulation: always complete
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float)(b->ms - a->ms); return cudaSuccess; }
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = nullptr; return cudaSuccess; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }
inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t*, void*) { return cudaErrorNotSupported; }
inline cudaError_t cudaIpcOpenMemHandle(void**, cudaIpcMemHandle_t, unsigned) { return cudaErrorNotSupported; }
inline cudaError_t cudaIpcCloseMemHandle(void*) { return cudaErrorNotSupported; }
template <class K>
inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, K, int, size_t) { *n = 1; return cudaSuccess; }
