/* cog_capi.cu — host side of the C ABI declared in include/cog_b200.h.
 *
 * Owns the device buffers of a handle (exported records in the reference's layouts
 * plus the private state), enqueues the kernels of cog_kernels.cuh on the handle's
 * stream and surfaces device-side failures.  There is no CPU implementation behind
 * this ABI: without a CUDA device every compute entry point fails with COG_ERR_CUDA.
 *
 * (tests/emu compiles this same file with g++ against its own <cuda_runtime.h> — a lock-step
 * warp emulator — to check the kernels' logic on machines without a GPU; that build is test
 * infrastructure and is never loaded by the package.  The only thing this file does for it is to
 * leave the kernel-launch statement overridable: COG_LAUNCH.)
 */
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <random>
#include <string>

#include "cog_b200.h"
#include "cog_kernels.cuh"

#ifndef COG_LAUNCH
#define COG_LAUNCH(kernel, grid, block, smem, stream, ...) \
  kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
/* programmatic dependent launch: the grid may become resident while the previous kernel of the stream
 * drains; the kernel itself waits (griddepcontrol.wait) before it touches memory */
template <class... P, class... A>
static void cog_launch_pdl(void (*kernel)(P...), unsigned grid, unsigned block, size_t smem, cudaStream_t stream, A... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}
#define COG_LAUNCH_PDL(kernel, grid, block, smem, stream, ...) cog_launch_pdl(kernel, grid, block, smem, stream, __VA_ARGS__)
#else
#define COG_LAUNCH_PDL COG_LAUNCH
#endif

static_assert(COG_OBS_SIZE == COGL_OBS_SIZE && COG_MASK_SIZE == COGL_MASK_SIZE &&
                  COG_ACTION_SIZE == COGL_ACTION_SIZE && COG_INFO_SIZE == COGL_INFO_SIZE,
              "C ABI and kernel layouts disagree");
static_assert(COG_STAT_COUNT == 12 && COG_STAT_RESETS == 10, "stat slots are hard-coded in the kernels");

namespace {

thread_local std::string g_error;
std::atomic<uint64_t> g_launches{0};

int fail(int code, const std::string &msg) {
  g_error = msg;
  return code;
}
int cuda_fail(cudaError_t e, const char *what) {
  return fail(COG_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CK(call)                                        \
  do {                                                  \
    cudaError_t e_ = (call);                            \
    if (e_ != cudaSuccess) return cuda_fail(e_, #call); \
  } while (0)

size_t scratch_bytes() { return sizeof(cog::BoardScratch) * cog::WARPS_PER_BLOCK; }
unsigned blocks_for(uint64_t n) { return (unsigned)((n + cog::WARPS_PER_BLOCK - 1) / cog::WARPS_PER_BLOCK); }
/* k_reset and k_step_single have block sizes of their own (COG_RESET_WARPS / COG_SINGLE_WARPS envs per block) */
unsigned blocks_of(uint64_t n, unsigned warps) { return (unsigned)((n + warps - 1) / warps); }
size_t scratch_of(unsigned warps) { return sizeof(cog::BoardScratch) * warps; }

}  // namespace

/* Which host-I/O path a handle uses for pinned arrays is measured, not assumed: on one GPU the
 * zero-copy kernels win (fewer bytes cross PCIe), with eight ranks hammering the same host
 * memory the staged DMA pipeline does.  Calls 0-1 warm the zero-copy path up, 2-4 time it,
 * 5 warms the pipeline up, 6-8 time it; from call 9 on the faster one is used.  Both paths
 * produce the same bytes.  COG_IO_ZEROCOPY=0 / 1 forces one of them. */
struct IoTune {
  int calls = 0, choice = -1; /* choice: 1 zero copy, 0 pipeline, -1 still measuring */
  double t_zero = 0, t_pipe = 0;
  bool use_zero() const { return choice >= 0 ? choice == 1 : calls < 5; }
  void record(bool zero, double seconds) {
    if (choice >= 0) return;
    if (zero && calls >= 2) t_zero += seconds;
    if (!zero && calls >= 6) t_pipe += seconds;
    if (++calls == 9) choice = t_zero <= t_pipe ? 1 : 0;
  }
};

struct cog_vec {
  uint64_t n = 0, offset = 0;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  cog::Buffers B{};
  cog::Params prm{100000, 4, 3, 0, 0}; /* cog_env::cog_env() defaults src/environment.cpp:10-15 */
  bool seeded = false;
  int sm_count = 1, fused_blocks_per_sm = 1, tail_blocks_per_sm = 1;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool timed = false;
  void *act_staging = nullptr; /* device copy of host actions for cog_vec_step_host */
  void *merged = nullptr;      /* COG_BUF_MERGED_MASKS, allocated on first use */
  void *compact = nullptr;     /* COG_BUF_COMPACT, allocated on first use */
  uint8_t *ring = nullptr;     /* rollout ring [n][ring_slots] compact records (cog_vec_rollout_attach) */
  uint32_t ring_slots = 0;
  uint64_t ring_steps = 0;     /* env-steps per env recorded since attach; next slot = ring_steps % ring_slots */
  cudaStream_t io[2] = {nullptr, nullptr}; /* copy/compute pipeline of the *_host_io entry points */
  cudaEvent_t io_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  IoTune tune;
  int *err_host = nullptr;     /* host address of B.err (mapped pinned memory) */
  cudaEvent_t ev_in = nullptr, ev_order = nullptr; /* input copy done / cross-stream ordering */
  void *pk_actions = nullptr, *pk_masks = nullptr; /* device staging of the packed host-I/O entry points */
  uint64_t host_steps = 0;     /* env-steps of single-step calls since the last stats reset (n per call) */
};

struct cog_sampler {
  uint64_t n = 0, offset = 0;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  uint32_t *srng = nullptr;
  uint8_t *actions = nullptr;
  void *staging = nullptr; /* device copy of host masks for cog_sampler_sample_host */
  void *pk_masks = nullptr, *pk_actions = nullptr; /* device staging of the packed host-I/O entry points */
  cudaEvent_t ev_in = nullptr, ev_order = nullptr;
  cudaStream_t io[2] = {nullptr, nullptr};
  cudaEvent_t io_ev[4] = {nullptr, nullptr, nullptr, nullptr};
  IoTune tune;
};

namespace {

/* Wait for the handle's stream, then report a device-side failure (board generation during a reset or an
 * auto-reset).  The flag lives in mapped pinned HOST memory (the kernels write it over PCIe only when
 * something fails), so every synchronising entry point can afford the check: no copy, no extra launch. */
int check_device_error(cog_vec *h) {
  CK(cudaStreamSynchronize(h->stream));
  volatile int *flag = h->err_host;
  if (flag[0] == 0) return COG_OK;
  const int err[2] = {flag[0], flag[1]};
  flag[0] = 0;
  flag[1] = 0;
  char msg[160];
  if (err[0] == COG_ERR_MAP_FAILURE)
    std::snprintf(msg, sizeof msg, "Failed to generate map in specified maximum number of attempts (env %d)", err[1]);
  else
    std::snprintf(msg, sizeof msg, "generated board does not fit the 48x48 observation (env %d)", err[1]);
  return fail(err[0], msg);
}

void time_begin(cog_vec *h) { cudaEventRecord(h->ev0, h->stream); }
void time_end(cog_vec *h) {
  cudaEventRecord(h->ev1, h->stream);
  h->timed = true;
}
/* Single steps are NOT bracketed by timing events unless COG_STEP_EVENTS=1: a timing event between two
 * kernels of a stream is a full serialisation point (measured: 44.0 us per sample+step of 65 536 envs with
 * the events, 38.9 us without, 36.5 us with programmatic dependent launch on top — profiles/r02_l_api_loop_events_pdl.txt). */
bool step_events() {
  static const bool on = std::getenv("COG_STEP_EVENTS") && std::atoi(std::getenv("COG_STEP_EVENTS")) != 0;
  return on;
}
/* programmatic dependent launch of the per-step kernels (COG_PDL=0 turns it off) */
bool use_pdl() {
  static const bool on = !(std::getenv("COG_PDL") && std::atoi(std::getenv("COG_PDL")) == 0);
  return on;
}

int launch_reset(cog_vec *h, uint32_t seed, int reseed) {
  CK(cudaSetDevice(h->device));
  COG_LAUNCH(cog::k_reset, blocks_of(h->n, COG_RESET_WARPS), COG_RESET_WARPS * 32, scratch_of(COG_RESET_WARPS), h->stream, h->B, h->prm,
             h->n, (uint64_t)seed + h->offset, reseed);
  g_launches++;
  CK(cudaGetLastError());
  return check_device_error(h);
}

/* ---- copy/compute pipeline of the *_host_io entry points -------------------------------
 * The batch is cut into chunks; chunk c runs on pipeline stream c & 1: H2D of its inputs,
 * its kernel, D2H of its results.  With two streams the H2D of one chunk, the kernel of
 * another and the D2H of a third overlap (both copy engines + SMs busy), so a host-buffer
 * call costs about max(H2D, D2H) over PCIe instead of their sum. */
struct Pipe {
  cudaStream_t *io;
  cudaEvent_t *ev; /* [0] pipeline -> main, [1] join, [2] fork, [3] main -> pipeline */
  cudaStream_t main;
};
/* Run `enqueue(io)` — which puts the chunks alternately on io[0] / io[1] — after everything
 * already on the handle's stream, and return when the host buffers are complete.
 * (Measured on the B200 box, PCIe 5: 8 MB move in 0.18 ms one way; every extra chunk adds
 * about 15 us of copy start-up latency in its stream, also when the sequence is replayed
 * from a CUDA graph — tried, no gain, removed — so two chunks is the sweet spot.) */
template <class F>
int pipe_run(Pipe &p, F enqueue) {
  for (int k = 0; k < 2; ++k)
    if (!p.io[k]) CK(cudaStreamCreateWithFlags(&p.io[k], cudaStreamNonBlocking));
  for (int k = 0; k < 4; ++k)
    if (!p.ev[k]) CK(cudaEventCreate(&p.ev[k]));
  CK(cudaEventRecord(p.ev[3], p.main));
  CK(cudaStreamWaitEvent(p.io[0], p.ev[3], 0));
  CK(cudaEventRecord(p.ev[2], p.io[0]));
  CK(cudaStreamWaitEvent(p.io[1], p.ev[2], 0));
  const int rc = enqueue(p.io);
  if (rc) return rc;
  CK(cudaEventRecord(p.ev[1], p.io[1]));
  CK(cudaStreamWaitEvent(p.io[0], p.ev[1], 0));
  CK(cudaEventRecord(p.ev[0], p.io[0]));
  CK(cudaStreamWaitEvent(p.main, p.ev[0], 0)); /* later work on the handle's stream comes after */
  CK(cudaStreamSynchronize(p.io[0]));          /* the host buffers are complete on return */
  return COG_OK;
}
/* Is `p` pinned host memory that kernels can address (cudaHostAlloc / cudaHostRegister under
 * unified addressing — what torch's pin_memory() hands out)?  Then the *_host_io entry points
 * skip the staging copies: the kernels read their inputs from, and write the bulky outputs to,
 * the caller's arrays directly, and only the bytes that are needed cross PCIe (3 of the 4
 * sectors of an ActionMask, 1 of the 2 of an ActionData).  COG_IO_ZEROCOPY=0 switches it off. */
int io_forced() { /* -1 measure, 0 pipeline, 1 zero copy */
  const char *e = std::getenv("COG_IO_ZEROCOPY");
  return !e ? -1 : e[0] == '0' ? 0 : 1;
}
double now_seconds() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
bool host_mapped(const void *p, void **dev) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  if (a.type != cudaMemoryTypeHost || !a.devicePointer) return false;
  *dev = a.devicePointer;
  return true;
}

uint64_t io_chunk(uint64_t n) {
  static const int chunks = [] { /* tunable: COG_IO_CHUNKS=1 disables the pipeline */
    const char *e = std::getenv("COG_IO_CHUNKS");
    const int v = e ? std::atoi(e) : 0;
    return v > 0 ? v : 2;
  }();
  uint64_t c = (n + chunks - 1) / chunks; /* two chunks by default ... */
  if (c < 4096) c = 4096;   /* ... unless that makes them too small to be worth a launch */
  return (c + cog::WARPS_PER_BLOCK - 1) / cog::WARPS_PER_BLOCK * cog::WARPS_PER_BLOCK;
}

}  // namespace

namespace {
/* one launch of the single-step kernel over the envs [i0, i1) on `stream` */
template <bool SAMPLED>
void launch_single(cog_vec *h, cudaStream_t stream, uint64_t i0, uint64_t i1, const void *actions,
                   uint32_t act_stride, void *masks2, void *pmasks, uint32_t *srng, uint8_t *actions_out) {
  if (use_pdl())
    COG_LAUNCH_PDL(cog::k_step_single<SAMPLED>, blocks_of(i1 - i0, COG_SINGLE_WARPS), COG_SINGLE_WARPS * 32,
                   scratch_of(COG_SINGLE_WARPS), stream, h->B,
                   h->prm, i0, i1, (const uint8_t *)actions, act_stride, (uint8_t *)masks2, (uint8_t *)pmasks, srng,
                   actions_out);
  else
    COG_LAUNCH(cog::k_step_single<SAMPLED>, blocks_of(i1 - i0, COG_SINGLE_WARPS), COG_SINGLE_WARPS * 32,
               scratch_of(COG_SINGLE_WARPS), stream, h->B,
               h->prm, i0, i1, (const uint8_t *)actions, act_stride, (uint8_t *)masks2, (uint8_t *)pmasks, srng,
               actions_out);
}

template <bool MERGED, bool ROLLOUT>
void launch_fused(cog_vec *h, cog_sampler *s, unsigned grid, unsigned tail_grid, uint32_t k_steps, const cog::Rollout &R) {
  auto hot = cog::k_step_fused<MERGED, ROLLOUT>;
  auto tail = cog::k_step_tail<MERGED, ROLLOUT>;
  if (ROLLOUT) { /* the ring-writing kernels have their own launch bounds: fill every SM exactly once */
    static int per_sm = 0;
    if (!per_sm && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hot, cog::BLOCK_THREADS, 0) != cudaSuccess)
      per_sm = 0;
    const unsigned full = blocks_for(h->n), resident = (unsigned)(h->sm_count * (per_sm > 0 ? per_sm : 1));
    grid = full < resident ? full : resident;
  }
  COG_LAUNCH(hot, grid, cog::BLOCK_THREADS, 0, h->stream, h->B, h->prm, h->n, s->srng, s->actions, k_steps, R);
  COG_LAUNCH(tail, tail_grid, cog::BLOCK_THREADS, scratch_bytes(), h->stream, h->B, h->prm, h->n, s->srng,
             s->actions, k_steps, R);
}
}  // namespace

extern "C" {

const char *cog_last_error(void) { return g_error.c_str(); }

int cog_device_count(int *count) {
  if (!count) return fail(COG_ERR_INVALID, "count is NULL");
  CK(cudaGetDeviceCount(count));
  return COG_OK;
}

uint64_t cog_kernel_launches(void) { return g_launches.load(); }

int cog_vec_create(uint64_t n_envs, uint64_t env_index_offset, int device, void *stream, cog_vec **out) {
  if (!out || n_envs == 0) return fail(COG_ERR_INVALID, "n_envs must be > 0 and out non-NULL");
  if (n_envs >= (1ull << 31)) return fail(COG_ERR_INVALID, "n_envs must be below 2^31 per handle (shard the job)");
  CK(cudaSetDevice(device));
  cog_vec *h = new (std::nothrow) cog_vec;
  if (!h) return fail(COG_ERR_INVALID, "out of host memory");
  /* from here on every failure path releases the handle and whatever it already owns */
#define CKH(call)                          \
  do {                                     \
    cudaError_t e_ = (call);               \
    if (e_ != cudaSuccess) {               \
      cog_vec_destroy(h);                  \
      return cuda_fail(e_, #call);         \
    }                                      \
  } while (0)
  h->n = n_envs;
  h->offset = env_index_offset;
  h->device = device;
  h->stream = (cudaStream_t)stream; /* NULL = the device's default stream */
  CKH(cudaEventCreate(&h->ev0));
  CKH(cudaEventCreate(&h->ev1));
  CKH(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device));
  const uint64_t n = n_envs;
  struct { void **p; size_t bytes; } bufs[] = {
      {(void **)&h->B.hidden, n * COG_HID_WORDS * sizeof(uint32_t)},
      {(void **)&h->B.grid, n * COGL_GRID_BYTES},
      {(void **)&h->B.obs, n * COG_OBS_SIZE},
      {(void **)&h->B.masks, n * COG_MASK_SIZE},
      {(void **)&h->B.rewards, n * 4 * sizeof(float)},
      {(void **)&h->B.dones, n},
      {(void **)&h->B.infos, n * COG_INFO_SIZE},
      {(void **)&h->B.agents, n},
      {(void **)&h->B.stats, COG_STAT_COUNT * sizeof(unsigned long long)},
      {(void **)&h->B.park, 4 * sizeof(unsigned)},
  };
  for (auto &b : bufs) {
    cudaError_t e = cudaMalloc(b.p, b.bytes);
    if (e == cudaSuccess) e = cudaMemsetAsync(*b.p, 0, b.bytes, h->stream);
    if (e != cudaSuccess) {
      cog_vec_destroy(h);
      return cuda_fail(e, "cudaMalloc");
    }
  }
  {
    cudaError_t e = cudaHostAlloc((void **)&h->err_host, 2 * sizeof(int), cudaHostAllocMapped);
    if (e == cudaSuccess) {
      h->err_host[0] = h->err_host[1] = 0;
      e = cudaHostGetDevicePointer((void **)&h->B.err, h->err_host, 0);
    }
    if (e != cudaSuccess) {
      cog_vec_destroy(h);
      return cuda_fail(e, "cudaHostAlloc (error flag)");
    }
  }
  CKH(cudaFuncSetAttribute(cog::k_step_tail<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scratch_bytes()));
  CKH(cudaFuncSetAttribute(cog::k_step_tail<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scratch_bytes()));
  CKH(cudaFuncSetAttribute(cog::k_step_tail<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scratch_bytes()));
  CKH(cudaFuncSetAttribute(cog::k_step_tail<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)scratch_bytes()));
  int per_sm = 0;
  CKH(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cog::k_step_fused<false, false>, cog::BLOCK_THREADS, 0));
  h->fused_blocks_per_sm = per_sm > 0 ? per_sm : 1;
  CKH(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cog::k_step_tail<false, false>, cog::BLOCK_THREADS, scratch_bytes()));
  h->tail_blocks_per_sm = per_sm > 0 ? per_sm : 1;
  COG_LAUNCH(cog::k_init, blocks_for(n), cog::BLOCK_THREADS, 0, h->stream, h->B, n);
  g_launches++;
  CKH(cudaGetLastError());
  CKH(cudaStreamSynchronize(h->stream));
  *out = h;
  return COG_OK;
#undef CKH
}

int cog_vec_destroy(cog_vec *h) {
  if (!h) return COG_OK;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  void *ptrs[] = {h->B.hidden, h->B.grid, h->B.obs, h->B.masks, h->B.rewards,
                  h->B.dones,  h->B.infos, h->B.agents, h->B.stats, h->B.park, h->act_staging, h->merged,
                  h->compact, h->ring, h->pk_actions, h->pk_masks};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  if (h->err_host) cudaFreeHost(h->err_host);
  if (h->ev_in) cudaEventDestroy(h->ev_in);
  if (h->ev_order) cudaEventDestroy(h->ev_order);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  for (cudaStream_t st : h->io)
    if (st) cudaStreamDestroy(st);
  for (cudaEvent_t e : h->io_ev)
    if (e) cudaEventDestroy(e);
  if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return COG_OK;
}

int cog_vec_num_envs(cog_vec *h, uint64_t *n) {
  if (!h || !n) return fail(COG_ERR_INVALID, "NULL argument");
  *n = h->n;
  return COG_OK;
}

int cog_vec_reset(cog_vec *h, uint32_t seed, uint8_t n_players, uint8_t n_pieces, int difficulty,
                  uint32_t max_steps, int render) {
  (void)render;
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  if (n_players < 1 || n_players > COG_MAX_PLAYERS) return fail(COG_ERR_INVALID, "n_players must be 1..4");
  if (difficulty < COG_EASY || difficulty > COG_HARD) return fail(COG_ERR_INVALID, "unknown difficulty");
  h->prm.max_steps = max_steps;
  h->prm.n_players = n_players;
  h->prm.n_pieces = n_pieces;
  h->prm.difficulty = (uint32_t)difficulty;
  h->seeded = true;
  return launch_reset(h, seed, 1);
}

int cog_vec_reset_default(cog_vec *h) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  if (!h->seeded) { /* cog_env() seeds from std::random_device (src/environment.cpp:10-15) */
    h->seeded = true;
    return launch_reset(h, std::random_device{}(), 1);
  }
  return launch_reset(h, 0, 0);
}

int cog_vec_step(cog_vec *h, const void *d_actions) {
  if (!h || !d_actions) return fail(COG_ERR_INVALID, "NULL argument");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "step() before reset()");
  CK(cudaSetDevice(h->device));
  if (step_events()) time_begin(h);
  launch_single<false>(h, h->stream, 0, h->n, d_actions, COG_ACTION_SIZE, nullptr, nullptr, nullptr, nullptr);
  if (step_events()) time_end(h);
  h->host_steps += h->n;
  g_launches++;
  CK(cudaGetLastError());
  return COG_OK;
}

int cog_vec_set_autoreset(cog_vec *h, int on) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  if (on) h->prm.flags &= ~(uint32_t)cog::COG_PRM_NO_AUTORESET;
  else h->prm.flags |= (uint32_t)cog::COG_PRM_NO_AUTORESET;
  return COG_OK;
}

int cog_vec_step_sampled(cog_vec *h, cog_sampler *s) {
  if (!h || !s) return fail(COG_ERR_INVALID, "NULL argument");
  if (s->n != h->n || s->device != h->device) return fail(COG_ERR_INVALID, "sampler and envs differ in size or device");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "step() before reset()");
  CK(cudaSetDevice(h->device));
  if (s->stream != h->stream) CK(cudaStreamSynchronize(s->stream));
  if (step_events()) time_begin(h);
  launch_single<true>(h, h->stream, 0, h->n, nullptr, COG_ACTION_SIZE, nullptr, nullptr, s->srng, s->actions);
  if (step_events()) time_end(h);
  h->host_steps += h->n;
  g_launches++;
  CK(cudaGetLastError());
  return COG_OK;
}

int cog_vec_buffer(cog_vec *h, int which, void **d_ptr, size_t *itemsize, size_t *count) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  void *p = nullptr;
  size_t item = 0;
  switch (which) {
  case COG_BUF_OBS: p = h->B.obs; item = COG_OBS_SIZE; break;
  case COG_BUF_SELECTED_MASKS: p = h->B.masks; item = COG_MASK_SIZE; break;
  case COG_BUF_REWARDS: p = h->B.rewards; item = 4 * sizeof(float); break;
  case COG_BUF_DONES: p = h->B.dones; item = 1; break;
  case COG_BUF_INFOS: p = h->B.infos; item = COG_INFO_SIZE; break;
  case COG_BUF_AGENT_SELECTION: p = h->B.agents; item = 1; break;
  case COG_BUF_MERGED_MASKS:
    if (!h->merged) {
      CK(cudaSetDevice(h->device));
      CK(cudaMalloc(&h->merged, h->n * COG_MASK_SIZE));
      CK(cudaMemsetAsync(h->merged, 0, h->n * COG_MASK_SIZE, h->stream));
    }
    p = h->merged; item = COG_MASK_SIZE; break;
  case COG_BUF_GRID: p = h->B.grid; item = COG_GRID_BYTES; break;
  case COG_BUF_COMPACT:
    if (!h->compact) {
      CK(cudaSetDevice(h->device));
      CK(cudaMalloc(&h->compact, h->n * COG_COMPACT_SIZE));
      CK(cudaMemsetAsync(h->compact, 0, h->n * COG_COMPACT_SIZE, h->stream));
    }
    p = h->compact; item = COG_COMPACT_SIZE; break;
  case COG_BUF_ROLLOUT:
    if (!h->ring) return fail(COG_ERR_INVALID, "no rollout ring attached");
    p = h->ring; item = COG_COMPACT_SIZE;
    if (d_ptr) *d_ptr = p;
    if (itemsize) *itemsize = item;
    if (count) *count = (size_t)h->n * h->ring_slots;
    return COG_OK;
  default: return fail(COG_ERR_INVALID, "unknown buffer id");
  }
  if (d_ptr) *d_ptr = p;
  if (itemsize) *itemsize = item;
  if (count) *count = (size_t)h->n;
  return COG_OK;
}

int cog_vec_read(cog_vec *h, int which, void *h_dst, size_t nbytes) {
  void *p;
  size_t item, count;
  int rc = cog_vec_buffer(h, which, &p, &item, &count);
  if (rc) return rc;
  if (!h_dst || nbytes > item * count) return fail(COG_ERR_INVALID, "bad destination / size");
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(h_dst, p, nbytes, cudaMemcpyDeviceToHost, h->stream));
  return check_device_error(h); /* synchronises */
}

int cog_sampler_create(uint64_t n_envs, uint64_t env_index_offset, uint32_t seed, int device, void *stream,
                       cog_sampler **out) {
  if (!out || n_envs == 0) return fail(COG_ERR_INVALID, "n_envs must be > 0 and out non-NULL");
  if (n_envs >= (1ull << 31)) return fail(COG_ERR_INVALID, "n_envs must be below 2^31 per handle (shard the job)");
  CK(cudaSetDevice(device));
  cog_sampler *s = new (std::nothrow) cog_sampler;
  if (!s) return fail(COG_ERR_INVALID, "out of host memory");
  s->n = n_envs;
  s->offset = env_index_offset;
  s->device = device;
  s->stream = (cudaStream_t)stream; /* NULL = the device's default stream */
  cudaError_t e = cudaMalloc((void **)&s->srng, n_envs * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMalloc((void **)&s->actions, n_envs * COG_ACTION_SIZE);
  if (e == cudaSuccess) e = cudaMemsetAsync(s->actions, 0, n_envs * COG_ACTION_SIZE, s->stream);
  if (e != cudaSuccess) {
    cog_sampler_destroy(s);
    return cuda_fail(e, "cudaMalloc");
  }
  COG_LAUNCH(cog::k_sampler_seed, (unsigned)((n_envs + 255) / 256), 256, 0, s->stream, s->srng, n_envs,
             (uint64_t)seed + env_index_offset);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(s->stream));
  *out = s;
  return COG_OK;
}

int cog_sampler_destroy(cog_sampler *s) {
  if (!s) return COG_OK;
  cudaSetDevice(s->device);
  cudaStreamSynchronize(s->stream);
  if (s->srng) cudaFree(s->srng);
  if (s->actions) cudaFree(s->actions);
  if (s->staging) cudaFree(s->staging);
  if (s->pk_masks) cudaFree(s->pk_masks);
  if (s->pk_actions) cudaFree(s->pk_actions);
  if (s->ev_in) cudaEventDestroy(s->ev_in);
  if (s->ev_order) cudaEventDestroy(s->ev_order);
  for (cudaStream_t st : s->io)
    if (st) cudaStreamDestroy(st);
  for (cudaEvent_t e : s->io_ev)
    if (e) cudaEventDestroy(e);
  if (s->own_stream && s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return COG_OK;
}

int cog_sampler_sample(cog_sampler *s, const void *d_masks) {
  if (!s || !d_masks) return fail(COG_ERR_INVALID, "NULL argument");
  CK(cudaSetDevice(s->device));
  COG_LAUNCH(cog::k_sample<false>, blocks_for(s->n), cog::BLOCK_THREADS, 0, s->stream, (const uint8_t *)d_masks,
             s->srng, s->actions, s->n, (uint8_t *)nullptr);
  g_launches++;
  CK(cudaGetLastError());
  return COG_OK;
}

int cog_sampler_sample_host(cog_sampler *s, const void *h_masks) {
  if (!s || !h_masks) return fail(COG_ERR_INVALID, "NULL argument");
  CK(cudaSetDevice(s->device));
  if (!s->staging) CK(cudaMalloc(&s->staging, s->n * COG_MASK_SIZE));
  CK(cudaMemcpyAsync(s->staging, h_masks, s->n * COG_MASK_SIZE, cudaMemcpyHostToDevice, s->stream));
  if (!s->ev_in) CK(cudaEventCreateWithFlags(&s->ev_in, cudaEventDisableTiming));
  CK(cudaEventRecord(s->ev_in, s->stream));
  const int rc = cog_sampler_sample(s, s->staging);
  /* the caller may rewrite its array as soon as we return (the reference's sample() is synchronous with
   * respect to its input): wait for the copy — not for the kernel, which stays asynchronous */
  CK(cudaEventSynchronize(s->ev_in));
  return rc;
}

int cog_sampler_actions(cog_sampler *s, void **d_ptr) {
  if (!s || !d_ptr) return fail(COG_ERR_INVALID, "NULL argument");
  *d_ptr = s->actions;
  return COG_OK;
}

int cog_sampler_read_actions(cog_sampler *s, void *h_dst, size_t nbytes) {
  if (!s || !h_dst || nbytes > s->n * COG_ACTION_SIZE) return fail(COG_ERR_INVALID, "bad destination / size");
  CK(cudaSetDevice(s->device));
  CK(cudaMemcpyAsync(h_dst, s->actions, nbytes, cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return COG_OK;
}

int cog_vec_step_host(cog_vec *h, const void *h_actions) {
  if (!h || !h_actions) return fail(COG_ERR_INVALID, "NULL argument");
  CK(cudaSetDevice(h->device));
  const size_t bytes = h->n * COG_ACTION_SIZE;
  if (!h->act_staging) CK(cudaMalloc(&h->act_staging, bytes));
  CK(cudaMemcpyAsync(h->act_staging, h_actions, bytes, cudaMemcpyHostToDevice, h->stream));
  if (!h->ev_in) CK(cudaEventCreateWithFlags(&h->ev_in, cudaEventDisableTiming));
  CK(cudaEventRecord(h->ev_in, h->stream));
  const int rc = cog_vec_step(h, h->act_staging);
  CK(cudaEventSynchronize(h->ev_in)); /* the caller's array is free again; the step itself stays asynchronous */
  return rc;
}

int cog_sampler_sample_host_io(cog_sampler *s, const void *h_masks, void *h_actions_out) {
  if (!s || !h_masks || !h_actions_out) return fail(COG_ERR_INVALID, "NULL argument");
  CK(cudaSetDevice(s->device));
  void *zm = nullptr, *za = nullptr;
  const int forced = io_forced();
  const bool eligible = forced != 0 && host_mapped(h_masks, &zm) && host_mapped(h_actions_out, &za);
  const bool measuring = eligible && forced < 0 && s->tune.choice < 0;
  const double t0 = measuring ? now_seconds() : 0.0;
  if (eligible && (forced == 1 || s->tune.use_zero())) {
    COG_LAUNCH(cog::k_sample<false>, blocks_for(s->n), cog::BLOCK_THREADS, 0, s->stream, (const uint8_t *)zm, s->srng,
               s->actions, s->n, (uint8_t *)za);
    g_launches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(s->stream)); /* the host array is complete on return */
    if (measuring) s->tune.record(true, now_seconds() - t0);
    return COG_OK;
  }
  if (!s->staging) CK(cudaMalloc(&s->staging, s->n * COG_MASK_SIZE));
  Pipe p{s->io, s->io_ev, s->stream};
  const uint64_t chunk = io_chunk(s->n);
  const int rc = pipe_run(p, [&](cudaStream_t *io) -> int {
    int k = 0;
    for (uint64_t i0 = 0; i0 < s->n; i0 += chunk, k ^= 1) {
      const uint64_t cnt = s->n - i0 < chunk ? s->n - i0 : chunk;
      uint8_t *d_m = (uint8_t *)s->staging + i0 * COG_MASK_SIZE;
      uint8_t *d_a = s->actions + i0 * COG_ACTION_SIZE;
      CK(cudaMemcpyAsync(d_m, (const uint8_t *)h_masks + i0 * COG_MASK_SIZE, cnt * COG_MASK_SIZE,
                         cudaMemcpyHostToDevice, io[k]));
      COG_LAUNCH(cog::k_sample<false>, blocks_for(cnt), cog::BLOCK_THREADS, 0, io[k], d_m, s->srng + i0, d_a, cnt,
                 (uint8_t *)nullptr);
      CK(cudaMemcpyAsync((uint8_t *)h_actions_out + i0 * COG_ACTION_SIZE, d_a, cnt * COG_ACTION_SIZE,
                         cudaMemcpyDeviceToHost, io[k]));
    }
    return COG_OK;
  });
  g_launches += (s->n + chunk - 1) / chunk;
  if (measuring && !rc) s->tune.record(false, now_seconds() - t0);
  return rc;
}

int cog_vec_step_host_io(cog_vec *h, const void *h_actions, void *h_masks_out, void *h_rewards_out,
                         void *h_dones_out, void *h_agents_out) {
  if (!h || !h_actions) return fail(COG_ERR_INVALID, "NULL argument");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "step() before reset()");
  CK(cudaSetDevice(h->device));
  void *za = nullptr, *zm = nullptr;
  const int forced = io_forced();
  const bool eligible = forced != 0 && host_mapped(h_actions, &za) && (!h_masks_out || host_mapped(h_masks_out, &zm));
  const bool measuring = eligible && forced < 0 && h->tune.choice < 0;
  const double t0 = measuring ? now_seconds() : 0.0;
  if (eligible && (forced == 1 || h->tune.use_zero())) {
    launch_single<false>(h, h->stream, 0, h->n, za, COG_ACTION_SIZE, zm, nullptr, nullptr, nullptr);
    h->host_steps += h->n;
    g_launches++;
    CK(cudaGetLastError());
    /* the small outputs by DMA (1.1 MB for 65 536 envs) */
    if (h_rewards_out)
      CK(cudaMemcpyAsync(h_rewards_out, h->B.rewards, h->n * 16, cudaMemcpyDeviceToHost, h->stream));
    if (h_dones_out) CK(cudaMemcpyAsync(h_dones_out, h->B.dones, h->n, cudaMemcpyDeviceToHost, h->stream));
    if (h_agents_out) CK(cudaMemcpyAsync(h_agents_out, h->B.agents, h->n, cudaMemcpyDeviceToHost, h->stream));
    const int rc = check_device_error(h); /* synchronises; a failed auto-reset is reported by the call that caused it */
    if (measuring) h->tune.record(true, now_seconds() - t0);
    return rc;
  }
  if (!h->act_staging) CK(cudaMalloc(&h->act_staging, h->n * COG_ACTION_SIZE));
  Pipe p{h->io, h->io_ev, h->stream};
  const uint64_t chunk = io_chunk(h->n);
  const int rc = pipe_run(p, [&](cudaStream_t *io) -> int {
    int k = 0;
    for (uint64_t i0 = 0; i0 < h->n; i0 += chunk, k ^= 1) {
      const uint64_t cnt = h->n - i0 < chunk ? h->n - i0 : chunk;
      cudaStream_t st = io[k];
      CK(cudaMemcpyAsync((uint8_t *)h->act_staging + i0 * COG_ACTION_SIZE,
                         (const uint8_t *)h_actions + i0 * COG_ACTION_SIZE, cnt * COG_ACTION_SIZE,
                         cudaMemcpyHostToDevice, st));
      launch_single<false>(h, st, i0, i0 + cnt, h->act_staging, COG_ACTION_SIZE, nullptr, nullptr, nullptr, nullptr);
      if (h_masks_out)
        CK(cudaMemcpyAsync((uint8_t *)h_masks_out + i0 * COG_MASK_SIZE, h->B.masks + i0 * COG_MASK_SIZE,
                           cnt * COG_MASK_SIZE, cudaMemcpyDeviceToHost, st));
      if (h_rewards_out)
        CK(cudaMemcpyAsync((uint8_t *)h_rewards_out + i0 * 16, (const uint8_t *)h->B.rewards + i0 * 16, cnt * 16,
                           cudaMemcpyDeviceToHost, st));
      if (h_dones_out)
        CK(cudaMemcpyAsync((uint8_t *)h_dones_out + i0, h->B.dones + i0, cnt, cudaMemcpyDeviceToHost, st));
      if (h_agents_out)
        CK(cudaMemcpyAsync((uint8_t *)h_agents_out + i0, h->B.agents + i0, cnt, cudaMemcpyDeviceToHost, st));
    }
    return COG_OK;
  });
  g_launches += (h->n + chunk - 1) / chunk;
  if (!rc) h->host_steps += h->n;
  if (measuring && !rc) h->tune.record(false, now_seconds() - t0);
  return rc ? rc : check_device_error(h);
}

/* ---- packed host I/O: 16-byte masks, 8-byte actions; staged through device memory by DMA (a 65 536-env
 * batch is 1 MB of masks and 0.5 MB of actions: two short copies beat 65 536 small PCIe reads) ---- */
int cog_sampler_sample_packed_host_io(cog_sampler *s, const void *h_pmasks, void *h_pactions_out) {
  if (!s || !h_pmasks || !h_pactions_out) return fail(COG_ERR_INVALID, "NULL argument");
  CK(cudaSetDevice(s->device));
  if (!s->pk_masks) CK(cudaMalloc(&s->pk_masks, s->n * COGL_PACKED_MASK_SIZE));
  if (!s->pk_actions) CK(cudaMalloc(&s->pk_actions, s->n * COGL_PACKED_ACTION_SIZE));
  CK(cudaMemcpyAsync(s->pk_masks, h_pmasks, s->n * COGL_PACKED_MASK_SIZE, cudaMemcpyHostToDevice, s->stream));
  COG_LAUNCH(cog::k_sample<true>, blocks_for(s->n), cog::BLOCK_THREADS, 0, s->stream, (const uint8_t *)s->pk_masks,
             s->srng, s->actions, s->n, (uint8_t *)s->pk_actions);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(h_pactions_out, s->pk_actions, s->n * COGL_PACKED_ACTION_SIZE, cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream)); /* the host array is complete on return */
  return COG_OK;
}

/* ---- the no-allocation variant of the call above (no counterpart in the reference; see cog_b200.h) ----
 * Staging for the packed masks (first part) and the packed actions (second part) of up to
 * COG_NOALLOC_MAX_ENVS envs.  One array per device that loads this library; every call on that device
 * shares it.  In HOST code the name below is only a handle for cudaGetSymbolAddress — never a pointer a
 * kernel may be given. */
namespace {
enum : size_t {
  NOALLOC_MASK_BYTES = (size_t)COG_NOALLOC_MAX_ENVS * COGL_PACKED_MASK_SIZE,
  NOALLOC_ACTION_BYTES = (size_t)COG_NOALLOC_MAX_ENVS * COGL_PACKED_ACTION_SIZE,
  NOALLOC_MAX_DEVICES = 64
};
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_MASK_BYTES + NOALLOC_ACTION_BYTES];
std::atomic<unsigned char *> g_noalloc_addr[NOALLOC_MAX_DEVICES];

/* device address of the array on `device` (which must be current); resolved once per device */
int noalloc_scratch(int device, unsigned char **out) {
  if (device < 0 || device >= (int)NOALLOC_MAX_DEVICES) return fail(COG_ERR_INVALID, "device ordinal out of range");
  unsigned char *p = g_noalloc_addr[device].load(std::memory_order_acquire);
  if (!p) {
    void *sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_noalloc_scratch);
    if (e != cudaSuccess || !sym) return cuda_fail(e, "cudaGetSymbolAddress(no-alloc scratch)");
    p = (unsigned char *)sym;
    g_noalloc_addr[device].store(p, std::memory_order_release); /* racing threads resolve the same address */
  }
  *out = p;
  return COG_OK;
}
}  // namespace

int cog_sampler_sample_packed_noalloc(cog_sampler *s, const void *h_pmasks, void *h_pactions_out) {
  if (!s || !h_pmasks || !h_pactions_out) return fail(COG_ERR_INVALID, "NULL argument");
  if (s->n > (uint64_t)COG_NOALLOC_MAX_ENVS)
    return fail(COG_ERR_INVALID, "cog_sampler_sample_packed_noalloc: sampler has " + std::to_string(s->n) +
                                     " envs, the static scratch holds at most COG_NOALLOC_MAX_ENVS = " +
                                     std::to_string((uint64_t)COG_NOALLOC_MAX_ENVS) +
                                     " (use cog_sampler_sample_packed_host_io, or shard the sampler)");
  CK(cudaSetDevice(s->device));
  unsigned char *scratch = nullptr;
  const int rc = noalloc_scratch(s->device, &scratch);
  if (rc) return rc;
  unsigned char *d_masks = scratch, *d_actions = scratch + NOALLOC_MASK_BYTES;
  CK(cudaMemcpyAsync(d_masks, h_pmasks, s->n * COGL_PACKED_MASK_SIZE, cudaMemcpyHostToDevice, s->stream));
  COG_LAUNCH(cog::k_sample<true>, blocks_for(s->n), cog::BLOCK_THREADS, 0, s->stream, (const uint8_t *)d_masks,
             s->srng, s->actions, s->n, (uint8_t *)d_actions);
  g_launches++;
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(h_pactions_out, d_actions, s->n * COGL_PACKED_ACTION_SIZE, cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream)); /* the host array is complete, and the shared scratch free, on return */
  return COG_OK;
}

int cog_vec_step_packed_host_io(cog_vec *h, const void *h_pactions, void *h_pmasks_out, void *h_rewards_out,
                                void *h_dones_out, void *h_agents_out) {
  if (!h || !h_pactions) return fail(COG_ERR_INVALID, "NULL argument");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "step() before reset()");
  CK(cudaSetDevice(h->device));
  if (!h->pk_actions) CK(cudaMalloc(&h->pk_actions, h->n * COGL_PACKED_ACTION_SIZE));
  if (!h->pk_masks) CK(cudaMalloc(&h->pk_masks, h->n * COGL_PACKED_MASK_SIZE));
  CK(cudaMemcpyAsync(h->pk_actions, h_pactions, h->n * COGL_PACKED_ACTION_SIZE, cudaMemcpyHostToDevice, h->stream));
  if (step_events()) time_begin(h);
  launch_single<false>(h, h->stream, 0, h->n, h->pk_actions, COGL_PACKED_ACTION_SIZE, nullptr,
                       h_pmasks_out ? h->pk_masks : nullptr, nullptr, nullptr);
  if (step_events()) time_end(h);
  h->host_steps += h->n;
  g_launches++;
  CK(cudaGetLastError());
  if (h_pmasks_out)
    CK(cudaMemcpyAsync(h_pmasks_out, h->pk_masks, h->n * COGL_PACKED_MASK_SIZE, cudaMemcpyDeviceToHost, h->stream));
  if (h_dones_out) CK(cudaMemcpyAsync(h_dones_out, h->B.dones, h->n, cudaMemcpyDeviceToHost, h->stream));
  if (h_agents_out) CK(cudaMemcpyAsync(h_agents_out, h->B.agents, h->n, cudaMemcpyDeviceToHost, h->stream));
  if (h_rewards_out) CK(cudaMemcpyAsync(h_rewards_out, h->B.rewards, h->n * 16, cudaMemcpyDeviceToHost, h->stream));
  return check_device_error(h); /* synchronises: the host arrays are complete on return */
}

int cog_vec_packed_masks(cog_vec *h, void *dst, int dst_is_host, int mask_mode) {
  if (!h || !dst) return fail(COG_ERR_INVALID, "NULL argument");
  if (mask_mode != COG_MASKS_SELECTED && mask_mode != COG_MASKS_MERGED) return fail(COG_ERR_INVALID, "unknown mask mode");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "masks before reset()");
  CK(cudaSetDevice(h->device));
  void *d = dst;
  if (dst_is_host) {
    if (!h->pk_masks) CK(cudaMalloc(&h->pk_masks, h->n * COGL_PACKED_MASK_SIZE));
    d = h->pk_masks;
  }
  COG_LAUNCH(cog::k_packed_masks, blocks_for(h->n), cog::BLOCK_THREADS, 0, h->stream, h->B, h->n, (uint8_t *)d,
             mask_mode == COG_MASKS_MERGED ? 1 : 0);
  g_launches++;
  CK(cudaGetLastError());
  if (dst_is_host) {
    CK(cudaMemcpyAsync(dst, d, h->n * COGL_PACKED_MASK_SIZE, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  return COG_OK;
}

/* ---- ordering against other streams (producers of the arrays handed to step / sample, consumers of the
 * exported buffers): an event on the earlier stream, a wait on the later one; no host synchronisation ---- */
namespace {
int order_streams(int device, cudaEvent_t *ev, cudaStream_t first, cudaStream_t then) {
  if (first == then) return COG_OK;
  CK(cudaSetDevice(device));
  if (!*ev) CK(cudaEventCreateWithFlags(ev, cudaEventDisableTiming));
  CK(cudaEventRecord(*ev, first));
  CK(cudaStreamWaitEvent(then, *ev, 0));
  return COG_OK;
}
}  // namespace
int cog_vec_wait_stream(cog_vec *h, void *producer) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  return order_streams(h->device, &h->ev_order, (cudaStream_t)producer, h->stream);
}
int cog_vec_release_to_stream(cog_vec *h, void *consumer) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  return order_streams(h->device, &h->ev_order, h->stream, (cudaStream_t)consumer);
}
int cog_sampler_wait_stream(cog_sampler *s, void *producer) {
  if (!s) return fail(COG_ERR_INVALID, "NULL handle");
  return order_streams(s->device, &s->ev_order, (cudaStream_t)producer, s->stream);
}
int cog_sampler_release_to_stream(cog_sampler *s, void *consumer) {
  if (!s) return fail(COG_ERR_INVALID, "NULL handle");
  return order_streams(s->device, &s->ev_order, s->stream, (cudaStream_t)consumer);
}
int cog_vec_stream(cog_vec *h, void **stream) {
  if (!h || !stream) return fail(COG_ERR_INVALID, "NULL argument");
  *stream = (void *)h->stream;
  return COG_OK;
}
int cog_sampler_stream(cog_sampler *s, void **stream) {
  if (!s || !stream) return fail(COG_ERR_INVALID, "NULL argument");
  *stream = (void *)s->stream;
  return COG_OK;
}

int cog_io_path(cog_vec *h, cog_sampler *s, int *vec_path, int *sampler_path) {
  const int forced = io_forced();
  if (vec_path) *vec_path = !h ? -1 : forced >= 0 ? forced : h->tune.choice;
  if (sampler_path) *sampler_path = !s ? -1 : forced >= 0 ? forced : s->tune.choice;
  return COG_OK;
}

int cog_vec_run_fused(cog_vec *h, cog_sampler *s, uint64_t k_steps) {
  return cog_vec_run_fused_ex(h, s, k_steps, COG_MASKS_SELECTED);
}

int cog_vec_compact_obs(cog_vec *h, void *d_out, int mask_mode) {
  if (!h) return fail(COG_ERR_INVALID, "NULL argument");
  if (mask_mode != COG_MASKS_SELECTED && mask_mode != COG_MASKS_MERGED) return fail(COG_ERR_INVALID, "unknown mask mode");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "compact observation before reset()");
  CK(cudaSetDevice(h->device));
  if (!d_out) {
    int rc = cog_vec_buffer(h, COG_BUF_COMPACT, &d_out, nullptr, nullptr);
    if (rc) return rc;
  }
  COG_LAUNCH(cog::k_compact_obs, blocks_for(h->n), cog::BLOCK_THREADS, 0, h->stream, h->B, h->prm, h->n,
             (uint8_t *)d_out, mask_mode == COG_MASKS_MERGED ? 1 : 0);
  g_launches++;
  CK(cudaGetLastError());
  return COG_OK;
}

int cog_vec_rollout_attach(cog_vec *h, uint32_t n_slots) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  if (h->ring) {
    CK(cudaFree(h->ring));
    h->ring = nullptr;
  }
  h->ring_slots = 0;
  h->ring_steps = 0;
  if (!n_slots) return COG_OK; /* detached */
  if ((uint64_t)n_slots * h->n > 0xffffffffull) return fail(COG_ERR_INVALID, "rollout ring: n_envs * n_slots must fit 32 bits");
  const size_t bytes = (size_t)n_slots * h->n * COG_COMPACT_SIZE;
  cudaError_t e = cudaMalloc((void **)&h->ring, bytes);
  if (e != cudaSuccess) {
    h->ring = nullptr;
    return cuda_fail(e, "cudaMalloc (rollout ring)");
  }
  CK(cudaMemsetAsync(h->ring, 0, bytes, h->stream));
  h->ring_slots = n_slots;
  return COG_OK;
}

int cog_vec_rollout_info(cog_vec *h, void **d_ring, uint32_t *n_slots, uint64_t *steps_recorded) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  if (d_ring) *d_ring = h->ring;
  if (n_slots) *n_slots = h->ring_slots;
  if (steps_recorded) *steps_recorded = h->ring_steps;
  return COG_OK;
}

int cog_vec_merged_masks(cog_vec *h, void *d_out) {
  if (!h) return fail(COG_ERR_INVALID, "NULL argument");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "merged masks before reset()");
  CK(cudaSetDevice(h->device));
  if (!d_out) {
    int rc = cog_vec_buffer(h, COG_BUF_MERGED_MASKS, &d_out, nullptr, nullptr);
    if (rc) return rc;
  }
  COG_LAUNCH(cog::k_merged_masks, blocks_for(h->n), cog::BLOCK_THREADS, 0, h->stream, h->B, h->n, (uint8_t *)d_out);
  g_launches++;
  CK(cudaGetLastError());
  return COG_OK;
}

int cog_vec_run_fused_ex(cog_vec *h, cog_sampler *s, uint64_t k_steps, int mask_mode) {
  if (!h || !s) return fail(COG_ERR_INVALID, "NULL argument");
  if (mask_mode != COG_MASKS_SELECTED && mask_mode != COG_MASKS_MERGED) return fail(COG_ERR_INVALID, "unknown mask mode");
  if (s->n != h->n || s->device != h->device) return fail(COG_ERR_INVALID, "sampler and envs differ in size or device");
  if (!h->seeded) return fail(COG_ERR_NOT_RESET, "run() before reset()");
  if (k_steps > 0xffffffffull) return fail(COG_ERR_INVALID, "k_steps must fit 32 bits");
  if (h->prm.flags & cog::COG_PRM_NO_AUTORESET)
    return fail(COG_ERR_INVALID, "the fused loop needs auto-reset (cog_vec_set_autoreset(h, 1))");
  CK(cudaSetDevice(h->device));
  if (s->stream != h->stream) CK(cudaStreamSynchronize(s->stream));
  /* persistent grid: every SM filled once, the warps draw their envs from a ticket counter (k_step_fused) */
  unsigned grid = blocks_for(h->n);
  const unsigned resident = (unsigned)(h->sm_count * h->fused_blocks_per_sm);
  if (grid > resident) grid = resident;
  cog::Rollout R{};
  if (h->ring) {
    R.ring = h->ring;
    R.T = h->ring_slots;
    R.slot0 = (uint32_t)(h->ring_steps % h->ring_slots);
    R.t0 = (uint32_t)h->ring_steps;
  }
  time_begin(h);
  if (k_steps) {
    /* the envs whose episode ended go to the tail kernel (exits at once when there are none);
     * it has its own resident grid */
    unsigned tail_grid = blocks_for(h->n);
    const unsigned tail_resident = (unsigned)(h->sm_count * h->tail_blocks_per_sm);
    if (tail_grid > tail_resident) tail_grid = tail_resident;
    const bool merged = mask_mode == COG_MASKS_MERGED;
    if (h->ring) {
      if (merged) launch_fused<true, true>(h, s, grid, tail_grid, (uint32_t)k_steps, R);
      else launch_fused<false, true>(h, s, grid, tail_grid, (uint32_t)k_steps, R);
      h->ring_steps += k_steps;
    } else {
      if (merged) launch_fused<true, false>(h, s, grid, tail_grid, (uint32_t)k_steps, R);
      else launch_fused<false, false>(h, s, grid, tail_grid, (uint32_t)k_steps, R);
    }
    g_launches += 2;
  }
  time_end(h);
  CK(cudaGetLastError());
  return COG_OK;
}

int cog_vec_stats(cog_vec *h, uint64_t *out, size_t n) {
  if (!h || !out || n > COG_STAT_COUNT) return fail(COG_ERR_INVALID, "bad argument");
  unsigned long long tmp[COG_STAT_COUNT];
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(tmp, h->B.stats, sizeof tmp, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  tmp[COG_STAT_ENV_STEPS] += h->host_steps; /* single steps are counted here, fused launches on the device */
  for (size_t i = 0; i < n; ++i) out[i] = tmp[i];
  return COG_OK;
}

int cog_vec_stats_reset(cog_vec *h) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  CK(cudaSetDevice(h->device));
  CK(cudaMemsetAsync(h->B.stats, 0, COG_STAT_COUNT * sizeof(unsigned long long), h->stream));
  h->host_steps = 0;
  return COG_OK;
}

int cog_sync(cog_vec *h) {
  if (!h) return fail(COG_ERR_INVALID, "NULL handle");
  CK(cudaSetDevice(h->device));
  return check_device_error(h);
}

int cog_vec_last_kernel_ms(cog_vec *h, float *ms) {
  if (!h || !ms) return fail(COG_ERR_INVALID, "NULL argument");
  if (!h->timed) return fail(COG_ERR_INVALID, "no kernel has been timed yet");
  CK(cudaSetDevice(h->device));
  CK(cudaEventSynchronize(h->ev1));
  CK(cudaEventElapsedTime(ms, h->ev0, h->ev1));
  return COG_OK;
}

}  /* extern "C" */
