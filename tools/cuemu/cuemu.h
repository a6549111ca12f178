// cuemu.h -- DEVELOPMENT AID ONLY: a tiny single-threaded CUDA *source-level* emulator used to
// debug the kernels' control flow on a machine without a GPU (this build container has none).
//
// It is NOT part of the product, is never linked into ebcc_b200/libh5z_ebcc.so and is not a CPU
// fallback: the product library calls the CUDA runtime unconditionally and fails loudly when no
// device is present.  The emulator compiles the very same .cu translation unit as host C++ with
// g++, running every CUDA thread of a CTA as a cooperative fiber (ucontext), CTAs one after the
// other.  __syncthreads()/__syncwarp()/shuffles/ballots are fiber rendezvous points, __shared__
// variables become function-local statics (legal because CTAs never overlap), and the subset of
// the runtime API the product uses (malloc/memcpy/memset/streams/events) maps onto libc.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <functional>

#define EBCC_CUEMU 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static
#define __align__(n) __attribute__((aligned(n)))

struct cuemu_dim3 { unsigned x, y, z; cuemu_dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
typedef cuemu_dim3 dim3;
extern cuemu_dim3 threadIdx, blockIdx, blockDim, gridDim;
static const int warpSize = 32;

struct uint2 { unsigned x, y; };
static inline uint2 make_uint2(unsigned a, unsigned b) { uint2 r = {a, b}; return r; }
struct float4 { float x, y, z, w; };
struct uint4 { unsigned x, y, z, w; };
static inline float4 make_float4(float a, float b, float c, float d) { float4 r = {a, b, c, d}; return r; }
static inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { uint4 r = {a, b, c, d}; return r; }

// ---- launch -------------------------------------------------------------------------------
void cuemu_launch(dim3 grid, dim3 block, size_t dyn_smem, const std::function<void()> &body);
extern unsigned char *cuemu_dyn_smem_ptr;
static unsigned long long g_ebcc_launches = 0;
#define EBCC_LAUNCH(kern, grid, block, smem, stream, ...) \
    ++g_ebcc_launches, cuemu_launch((grid), (block), (smem), [=]() { kern(__VA_ARGS__); })
#define EBCC_DYN_SMEM(type, name) type *name = (type *)cuemu_dyn_smem_ptr

// ---- synchronisation ----------------------------------------------------------------------
void __syncthreads();
int __syncthreads_or(int pred);
int __syncthreads_count(int pred);
int __syncthreads_and(int pred);
void __syncwarp(unsigned mask = 0xffffffffu);
unsigned __ballot_sync(unsigned mask, int pred);
int __any_sync(unsigned mask, int pred);
int __all_sync(unsigned mask, int pred);
uint64_t cuemu_shfl(uint64_t v, int src_lane);        // generic exchange (all 32 lanes participate)
uint64_t cuemu_shfl_partial(unsigned mask, uint64_t v, int src_lane);
template <typename T> static inline T __shfl_sync(unsigned mask, T v, int src, int width = 32)
{
    uint64_t raw = 0; memcpy(&raw, &v, sizeof(T));
    int lane = (int)(threadIdx.x & 31);
    int s = (lane & ~(width - 1)) | (src & (width - 1));
    raw = cuemu_shfl_partial(mask, raw, s);
    T r; memcpy(&r, &raw, sizeof(T)); return r;
}
template <typename T> static inline T __shfl_down_sync(unsigned mask, T v, unsigned d, int width = 32)
{
    int lane = (int)(threadIdx.x & 31);
    int s = ((lane & (width - 1)) + (int)d < width) ? lane + (int)d : lane;
    return __shfl_sync(mask, v, s, 32);
}
template <typename T> static inline T __shfl_up_sync(unsigned mask, T v, unsigned d, int width = 32)
{
    int lane = (int)(threadIdx.x & 31);
    int s = ((lane & (width - 1)) >= (int)d) ? lane - (int)d : lane;
    return __shfl_sync(mask, v, s, 32);
}
template <typename T> static inline T __shfl_xor_sync(unsigned mask, T v, int m, int width = 32)
{
    int lane = (int)(threadIdx.x & 31);
    return __shfl_sync(mask, v, lane ^ m, 32);
}
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline void __nanosleep(unsigned) {}

// ---- atomics (single OS thread: plain read-modify-write is atomic) ----------------------------
template <typename T> static inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; return o; }
template <typename T> static inline T atomicMin(T *p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <typename T> static inline T atomicMax(T *p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <typename T> static inline T atomicOr(T *p, T v) { T o = *p; *p = o | v; return o; }
template <typename T> static inline T atomicAnd(T *p, T v) { T o = *p; *p = o & v; return o; }
template <typename T> static inline T atomicExch(T *p, T v) { T o = *p; *p = v; return o; }
template <typename T> static inline T atomicCAS(T *p, T c, T v) { T o = *p; if (o == c) *p = v; return o; }

// ---- intrinsics -----------------------------------------------------------------------------
// the emulator TU is compiled with -ffp-contract=off, so plain operators are IEEE RN without FMA
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __ddiv_rn(double a, double b) { return a / b; }
static inline int __float2int_rn(float a)      // saturating like cvt.rni.s32.f32; NaN -> 0
{
    if (!(a == a)) return 0;
    if (a >= 2147483648.0f) return 2147483647;
    if (a <= -2147483648.0f) return -2147483647 - 1;
    return (int)lrintf(a);
}
static inline long long __double2ll_rn(double a) { return llrint(a); }
static inline double __longlong_as_double(long long a) { double d; memcpy(&d, &a, 8); return d; }
static inline int __float2int_rz(float a) { return (int)a; }
static inline unsigned __float2uint_rz(float a) { return a <= 0 ? 0u : (unsigned)a; }
static inline float __int2float_rn(int a) { return (float)a; }
static inline float __uint2float_rn(unsigned a) { return (float)a; }
static inline unsigned __float_as_uint(float f) { unsigned u; memcpy(&u, &f, 4); return u; }
static inline int __float_as_int(float f) { int u; memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(unsigned u) { float f; memcpy(&f, &u, 4); return f; }
static inline float __int_as_float(int u) { float f; memcpy(&f, &u, 4); return f; }
static inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
static inline int __clzll(long long x) { return x == 0 ? 64 : __builtin_clzll((unsigned long long)x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
static inline unsigned __brev(unsigned x) { unsigned r = 0; for (int i = 0; i < 32; ++i) if (x >> i & 1) r |= 1u << (31 - i); return r; }
static inline unsigned __byte_perm(unsigned a, unsigned b, unsigned sel)
{
    unsigned long long v = ((unsigned long long)b << 32) | a; unsigned r = 0;
    for (int i = 0; i < 4; ++i) r |= (unsigned)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
    return r;
}
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned s) { s &= 31; return s ? (hi << s) | (lo >> (32 - s)) : hi; }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s) { s &= 31; return s ? (lo >> s) | (hi << (32 - s)) : lo; }
template <typename T> static inline T __ldg(const T *p) { return *p; }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
static inline unsigned umin(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned umax(unsigned a, unsigned b) { return a > b ? a : b; }
static inline float fminf_(float a, float b) { return fminf(a, b); }

// ---- runtime API subset -----------------------------------------------------------------------
typedef int cudaError_t;
typedef intptr_t cudaStream_t;
typedef struct cuemu_event { double t; } *cudaEvent_t;
enum { cudaSuccess = 0, cudaErrorNotReady = 600 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum { cudaHostAllocDefault = 0, cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2, cudaEventBlockingSync = 1 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize, cudaFuncAttributePreferredSharedMemoryCarveout };
enum { cudaSharedmemCarveoutMaxShared = 100 };
static inline const char *cudaGetErrorString(cudaError_t) { return "cuemu"; }
static inline cudaError_t cudaGetLastError() { return 0; }
static inline cudaError_t cudaPeekAtLastError() { return 0; }
static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = malloc(n ? n : 1); return *p ? 0 : 2; }
template <typename T> static inline cudaError_t cudaMalloc(T **p, size_t n) { return cudaMalloc((void **)p, n); }
static inline cudaError_t cudaFree(void *p) { free(p); return 0; }
static inline cudaError_t cudaMallocHost(void **p, size_t n) { *p = malloc(n ? n : 1); return 0; }
template <typename T> static inline cudaError_t cudaMallocHost(T **p, size_t n) { return cudaMallocHost((void **)p, n); }
static inline cudaError_t cudaHostAlloc(void **p, size_t n, unsigned) { *p = malloc(n ? n : 1); return 0; }
static inline cudaError_t cudaFreeHost(void *p) { free(p); return 0; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = 0) { memmove(d, s, n); return 0; }
static inline cudaError_t cudaMemcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, cudaMemcpyKind, cudaStream_t = 0)
{
    for (size_t r = 0; r < h; ++r) memmove((char *)d + r * dp, (const char *)s + r * sp, w);
    return 0;
}
static inline cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t = 0) { memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamCreate(cudaStream_t *s) { *s = 0; return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = 0; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaSetDevice(int) { return 0; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return 0; }
static inline cudaError_t cudaGetDeviceCount(int *d) { *d = 1; return 0; }
static inline cudaError_t cudaDeviceGetStreamPriorityRange(int *l, int *g) { *l = 0; *g = 0; return 0; }
static inline cudaError_t cudaStreamCreateWithPriority(cudaStream_t *s, unsigned, int) { *s = 0; return 0; }
cudaError_t cudaEventCreate(cudaEvent_t *e);
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { return cudaEventCreate(e); }
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s = 0);
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
static inline cudaError_t cudaEventQuery(cudaEvent_t) { return 0; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return 0; }
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b);
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { free(e); return 0; }
template <typename F> static inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return 0; }
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &sym) { *p = (void *)&sym; return 0; }
template <typename T> static inline cudaError_t cudaMemcpyToSymbol(T &sym, const void *src, size_t n, size_t off = 0, cudaMemcpyKind = cudaMemcpyHostToDevice)
{ memcpy((char *)&sym + off, src, n); return 0; }
