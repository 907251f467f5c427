/* cuda_emu.h — TEST INFRASTRUCTURE: a lock-step warp emulator for machines without a GPU.
 *
 * Compiles the package's CUDA sources (gym_eldorado_b200/csrc/*.cu[h]) unchanged with
 * g++ so that `pytest -m "not gpu"` can check the kernels' game logic against the
 * oracle.  Each of a warp's 32 lanes runs the kernel body as a fiber; a lane runs
 * until it reaches a warp collective (__shfl_sync, __ballot_sync, __syncwarp, ...),
 * and when all 32 lanes have arrived at the same collective the scheduler resolves it
 * and resumes them.  A lane that exits while others wait, or lanes that meet at
 * different collectives, abort the run — divergence bugs show up as failures.
 *
 * This is NOT a CPU backend of the product: nothing under gym_eldorado_b200/ loads the
 * emulated library, it is single threaded and ~1000x slower than the GPU, and it says
 * nothing about memory-model or performance behaviour.  Only tests/ builds and loads it.
 * Include in exactly one translation unit.
 */
#ifndef CUDA_EMU_H
#define CUDA_EMU_H
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>

#define __device__
#define __global__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __shared__
#define __constant__
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))

using std::max;
using std::min;

struct uint4 { uint32_t x, y, z, w; };
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
struct uint2 { uint32_t x, y; };
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{x, y}; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
struct emu_dim3 { unsigned x = 1, y = 1, z = 1; };
static emu_dim3 threadIdx, blockIdx, blockDim, gridDim;

/* ---- context switch (x86-64 SysV): callee-saved registers + stack pointer ---- */
extern "C" void cuda_emu_switch(void **save_sp, void *next_sp);
__asm__(
    ".text\n"
    ".globl cuda_emu_switch\n"
    ".type cuda_emu_switch,@function\n"
    "cuda_emu_switch:\n"
    "  pushq %rbp\n  pushq %rbx\n  pushq %r12\n  pushq %r13\n  pushq %r14\n  pushq %r15\n"
    "  movq %rsp, (%rdi)\n"
    "  movq %rsi, %rsp\n"
    "  popq %r15\n  popq %r14\n  popq %r13\n  popq %r12\n  popq %rbx\n  popq %rbp\n"
    "  ret\n"
    ".size cuda_emu_switch,.-cuda_emu_switch\n");

namespace cuda_emu {

enum Op { OP_NONE = 0, OP_SYNC, OP_SHFL, OP_SHFL_UP, OP_SHFL_XOR, OP_BALLOT, OP_ANY, OP_ALL,
          OP_RED_ADD, OP_RED_MIN_S, OP_RED_MAX_S, OP_RED_MIN_U, OP_RED_MAX_U, OP_RED_OR };

struct Fiber { void *sp = nullptr; bool done = true; char *stack = nullptr; int op = OP_NONE; };
constexpr size_t STACK_BYTES = 512 * 1024;
static Fiber g_fiber[32];
static void *g_sched_sp = nullptr;
static int g_lane = 0;
static const std::function<void()> *g_body = nullptr;
static uint32_t g_in[32], g_arg[32], g_out[32];

[[noreturn]] static void die(const char *msg) {
  std::fprintf(stderr, "cuda_emu: %s\n", msg);
  std::abort();
}

static void fiber_entry() {
  (*g_body)();
  g_fiber[g_lane].done = true;
  g_fiber[g_lane].op = OP_NONE;
  cuda_emu_switch(&g_fiber[g_lane].sp, g_sched_sp);
  die("resumed a finished lane");
}

static inline uint32_t collective(int op, uint32_t in, uint32_t arg) {
  const int l = g_lane;
  g_in[l] = in;
  g_arg[l] = arg;
  g_fiber[l].op = op;
  cuda_emu_switch(&g_fiber[l].sp, g_sched_sp);
  return g_out[l];
}

static void resolve(int op) {
  uint32_t acc = 0;
  switch (op) {
  case OP_SYNC: break;
  case OP_SHFL: for (int l = 0; l < 32; ++l) g_out[l] = g_in[g_arg[l] & 31]; break;
  case OP_SHFL_UP: for (int l = 0; l < 32; ++l) g_out[l] = l >= (int)g_arg[l] ? g_in[l - g_arg[l]] : g_in[l]; break;
  case OP_SHFL_XOR: for (int l = 0; l < 32; ++l) g_out[l] = g_in[(l ^ g_arg[l]) & 31]; break;
  case OP_BALLOT: case OP_ANY: case OP_ALL:
    for (int l = 0; l < 32; ++l) acc |= (g_in[l] ? 1u : 0u) << l;
    if (op == OP_ANY) acc = acc != 0;
    if (op == OP_ALL) acc = acc == 0xffffffffu;
    for (int l = 0; l < 32; ++l) g_out[l] = acc;
    break;
  case OP_RED_ADD: for (int l = 0; l < 32; ++l) acc += g_in[l];
    for (int l = 0; l < 32; ++l) g_out[l] = acc; break;
  case OP_RED_OR: for (int l = 0; l < 32; ++l) acc |= g_in[l];
    for (int l = 0; l < 32; ++l) g_out[l] = acc; break;
  case OP_RED_MIN_S: case OP_RED_MAX_S: {
    int32_t v = (int32_t)g_in[0];
    for (int l = 1; l < 32; ++l) v = op == OP_RED_MIN_S ? std::min(v, (int32_t)g_in[l]) : std::max(v, (int32_t)g_in[l]);
    for (int l = 0; l < 32; ++l) g_out[l] = (uint32_t)v;
    break;
  }
  case OP_RED_MIN_U: case OP_RED_MAX_U: {
    uint32_t v = g_in[0];
    for (int l = 1; l < 32; ++l) v = op == OP_RED_MIN_U ? std::min(v, g_in[l]) : std::max(v, g_in[l]);
    for (int l = 0; l < 32; ++l) g_out[l] = v;
    break;
  }
  default: die("unknown collective");
  }
}

static void run_warp(unsigned block, unsigned warp, const std::function<void()> &body) {
  g_body = &body;
  for (int l = 0; l < 32; ++l) {
    Fiber &f = g_fiber[l];
    if (!f.stack) f.stack = (char *)std::malloc(STACK_BYTES);
    void **sp = (void **)(((uintptr_t)(f.stack + STACK_BYTES)) & ~(uintptr_t)15);
    *--sp = nullptr;                /* keeps rsp = 8 mod 16 at fiber_entry */
    *--sp = (void *)&fiber_entry;   /* return address consumed by `ret` */
    for (int k = 0; k < 6; ++k) *--sp = nullptr;
    f.sp = sp;
    f.done = false;
    f.op = OP_NONE;
  }
  for (;;) {
    int alive = 0;
    for (int l = 0; l < 32; ++l) {
      if (g_fiber[l].done) continue;
      g_lane = l;
      blockIdx.x = block;
      threadIdx.x = warp * 32 + (unsigned)l;
      cuda_emu_switch(&g_sched_sp, g_fiber[l].sp);
      if (!g_fiber[l].done) ++alive;
    }
    if (!alive) return;
    int op = OP_NONE;
    for (int l = 0; l < 32; ++l) {
      if (g_fiber[l].done) die("a lane exited while others wait at a full-mask collective");
      if (op == OP_NONE) op = g_fiber[l].op;
      else if (op != g_fiber[l].op) die("lanes of one warp met at different collectives");
    }
    resolve(op);
  }
}

template <class F>
static void launch(unsigned grid, unsigned block, size_t /*smem*/, F &&f) {
  const std::function<void()> body(f);
  if (block % 32) die("block size must be a multiple of 32");
  gridDim.x = grid;
  blockDim.x = block;
  for (unsigned b = 0; b < grid; ++b)
    for (unsigned w = 0; w < block / 32; ++w) run_warp(b, w, body);
}

}  // namespace cuda_emu

/* ---- warp intrinsics ---- */
static inline void __syncwarp(unsigned = 0xffffffffu) { cuda_emu::collective(cuda_emu::OP_SYNC, 0, 0); }
template <class T> static inline T __shfl_sync(unsigned, T v, int src) {
  static_assert(sizeof(T) == 4, "32-bit shuffles only");
  uint32_t u; std::memcpy(&u, &v, 4);
  u = cuda_emu::collective(cuda_emu::OP_SHFL, u, (uint32_t)src);
  std::memcpy(&v, &u, 4); return v;
}
template <class T> static inline T __shfl_up_sync(unsigned, T v, unsigned d) {
  static_assert(sizeof(T) == 4, "32-bit shuffles only");
  uint32_t u; std::memcpy(&u, &v, 4);
  u = cuda_emu::collective(cuda_emu::OP_SHFL_UP, u, d);
  std::memcpy(&v, &u, 4); return v;
}
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m) {
  static_assert(sizeof(T) == 4, "32-bit shuffles only");
  uint32_t u; std::memcpy(&u, &v, 4);
  u = cuda_emu::collective(cuda_emu::OP_SHFL_XOR, u, (uint32_t)m);
  std::memcpy(&v, &u, 4); return v;
}
static inline unsigned __ballot_sync(unsigned, int pred) { return cuda_emu::collective(cuda_emu::OP_BALLOT, pred != 0, 0); }
static inline int __any_sync(unsigned, int pred) { return (int)cuda_emu::collective(cuda_emu::OP_ANY, pred != 0, 0); }
static inline int __all_sync(unsigned, int pred) { return (int)cuda_emu::collective(cuda_emu::OP_ALL, pred != 0, 0); }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) { return cuda_emu::collective(cuda_emu::OP_RED_ADD, v, 0); }
static inline int __reduce_add_sync(unsigned, int v) { return (int)cuda_emu::collective(cuda_emu::OP_RED_ADD, (uint32_t)v, 0); }
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return cuda_emu::collective(cuda_emu::OP_RED_OR, v, 0); }
static inline int __reduce_min_sync(unsigned, int v) { return (int)cuda_emu::collective(cuda_emu::OP_RED_MIN_S, (uint32_t)v, 0); }
static inline int __reduce_max_sync(unsigned, int v) { return (int)cuda_emu::collective(cuda_emu::OP_RED_MAX_S, (uint32_t)v, 0); }
static inline unsigned __reduce_min_sync(unsigned, unsigned v) { return cuda_emu::collective(cuda_emu::OP_RED_MIN_U, v, 0); }
static inline unsigned __reduce_max_sync(unsigned, unsigned v) { return cuda_emu::collective(cuda_emu::OP_RED_MAX_U, v, 0); }

static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline unsigned __fns(unsigned mask, unsigned base, int offset) {
  if (offset <= 0 || base != 0) cuda_emu::die("__fns: only base 0 / positive offsets are emulated");
  for (unsigned b = 0; b < 32; ++b)
    if ((mask >> b) & 1u)
      if (--offset == 0) return b;
  return 0xffffffffu;
}
static inline unsigned __byte_perm(unsigned a, unsigned b, unsigned sel) {
  const uint64_t both = ((uint64_t)b << 32) | a;
  unsigned r = 0;
  for (int i = 0; i < 4; ++i) {
    const unsigned n = (sel >> (4 * i)) & 0xf;
    unsigned byte = (unsigned)(both >> (8 * (n & 7))) & 0xffu;
    if (n & 8) byte = (byte & 0x80u) ? 0xffu : 0u; /* sign-replicate mode */
    r |= byte << (8 * i);
  }
  return r;
}
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((uint64_t)a * b) >> 32); }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s) {
  return (unsigned)(((((uint64_t)hi) << 32) | lo) >> (s & 31u));
}

template <class T> static inline T __ldcg(const T *p) { return *p; } /* cache hints mean nothing here */
static inline void __threadfence() { __asm__ __volatile__("" ::: "memory"); } /* warps run one after another */

/* lanes run one at a time, so plain read-modify-write is atomic */
template <class T> static inline T atomicOr(T *p, T v) { T o = *p; *p = o | v; return o; }
template <class T> static inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; return o; }
template <class T> static inline T atomicCAS(T *p, T cmp, T v) { T o = *p; if (o == cmp) *p = v; return o; }

/* ---- the slice of the CUDA runtime the host code uses ---- */
typedef int cudaError_t;
typedef void *cudaStream_t;
struct cuda_emu_event { std::chrono::steady_clock::time_point t; };
typedef cuda_emu_event *cudaEvent_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
enum { cudaDevAttrMultiProcessorCount = 16, cudaStreamNonBlocking = 1 };
static inline const char *cudaGetErrorString(cudaError_t) { return "emulated"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaSetDevice(int d) { return d == 0 ? cudaSuccess : 1; }
static inline cudaError_t cudaGetDeviceCount(int *c) { *c = 1; return cudaSuccess; }
static inline cudaError_t cudaDeviceGetAttribute(int *v, int, int) { *v = 1; return cudaSuccess; }
static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? cudaSuccess : 2; }
static inline cudaError_t cudaFree(void *p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void *p, int v, size_t n, cudaStream_t) { std::memset(p, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t) { std::memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = nullptr; return cudaSuccess; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
/* the product's kernel-launch statement (csrc/cog_capi.cu: COG_LAUNCH), overridden: blocks and warps run one
 * after the other on the calling thread */
#define COG_LAUNCH(kernel, grid, block, smem, stream, ...) \
  cuda_emu::launch((grid), (block), (smem), [&]() { kernel(__VA_ARGS__); })
/* "device" memory is host memory: every pointer is device addressable */
enum cudaMemoryType { cudaMemoryTypeUnregistered = 0, cudaMemoryTypeHost = 1, cudaMemoryTypeDevice = 2 };
struct cudaPointerAttributes { cudaMemoryType type; int device; void *devicePointer; void *hostPointer; };
static inline cudaError_t cudaPointerGetAttributes(cudaPointerAttributes *a, const void *p) {
  a->type = cudaMemoryTypeHost; a->device = 0; a->devicePointer = a->hostPointer = const_cast<void *>(p);
  return cudaSuccess;
}
/* a `__device__` array is plain host memory here: its address is the symbol's address */
template <class T> static inline cudaError_t cudaGetSymbolAddress(void **p, const T &symbol) { *p = const_cast<void *>(static_cast<const void *>(&symbol)); return cudaSuccess; }
enum { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return cudaSuccess; }
template <class F> static inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int *n, F, int, size_t) { *n = 1; return cudaSuccess; }
enum { cudaHostAllocMapped = 2, cudaEventDisableTiming = 2 };
static inline cudaError_t cudaHostAlloc(void **p, size_t n, unsigned) { *p = std::malloc(n ? n : 1); return *p ? cudaSuccess : 2; }
static inline cudaError_t cudaFreeHost(void *p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaHostGetDevicePointer(void **d, void *h, unsigned) { *d = h; return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new cuda_emu_event; return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { *e = new cuda_emu_event; return cudaSuccess; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) { e->t = std::chrono::steady_clock::now(); return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) {
  *ms = std::chrono::duration<float, std::milli>(b->t - a->t).count();
  return cudaSuccess;
}
#endif
