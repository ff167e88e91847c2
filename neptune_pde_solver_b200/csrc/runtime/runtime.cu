// neptune_rt_* — the B200 runtime behind the `neptuneir-to-b200` pipeline (include/neptune_rt.h).
//
// Plays the role lib/Runtime/PETSc/NeptunePETScRuntime.cpp plays for `_neptune_rt_runtime_*`
// in the reference, for the one op this project accelerates: `neptune_ir.apply`, which the
// reference lowers to malloc + memref.copy + a scalar scf.for nest
// (lib/Passes/DataflowLowering.cpp:258-448).  Fields live in HBM as identity-layout row-major
// boxes (NeptuneIRTypeConverter, DataflowLowering.cpp:41-71); a sweep is one launch of a
// hand-written sm_100a kernel instantiated for the apply's body (kernels/*.cuh).
// There is no CPU path anywhere in this file: without a CUDA device every entry point fails.
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <nvtx3/nvToolsExt.h>   // header-only: ranges show up when a profiler injects, cost nothing otherwise

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <functional>
#include <initializer_list>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <vector>

#include "frontend/ir.h"
#include "kernels/registry.h"
#include "neptune_rt.h"

using neptune::BodyEntry;
using neptune::BodyInfo;
using neptune::StencilDesc;
using neptune::SweepArgs;
using neptune::Variant;

// ---------------------------------------------------------------------------------------------
// NCCL is bound at run time: inside a torch process the library torch already loaded is reused
// (one NCCL per process), elsewhere the system libnccl.so.2 is opened.
// ---------------------------------------------------------------------------------------------
namespace {
struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*CommGetAsyncError)(ncclComm_t, ncclResult_t *) = nullptr;   // optional
  bool ok = false;
};

NcclApi &nccl() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api;
  tried = true;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char *n : names) {
    api.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (api.handle) break;
  }
  if (!api.handle) {
    // a symbol already in the process (torch links its own copy)?
    if (dlsym(RTLD_DEFAULT, "ncclGetUniqueId")) api.handle = RTLD_DEFAULT;
  }
  if (!api.handle)
    for (const char *n : names) {
      api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
  if (!api.handle) return api;
#define NB_SYM(field, sym) api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, sym))
  NB_SYM(GetUniqueId, "ncclGetUniqueId");
  NB_SYM(CommInitRank, "ncclCommInitRank");
  NB_SYM(CommDestroy, "ncclCommDestroy");
  NB_SYM(Send, "ncclSend");
  NB_SYM(Recv, "ncclRecv");
  NB_SYM(AllReduce, "ncclAllReduce");
  NB_SYM(GroupStart, "ncclGroupStart");
  NB_SYM(GroupEnd, "ncclGroupEnd");
  NB_SYM(GetErrorString, "ncclGetErrorString");
  NB_SYM(CommGetAsyncError, "ncclCommGetAsyncError");
#undef NB_SYM
  api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.Send && api.Recv && api.AllReduce && api.GroupStart &&
           api.GroupEnd && api.GetErrorString;
  return api;
}

thread_local std::string g_last_error;
std::set<neptune_rt_ctx *> &live_ctxs() {
  static std::set<neptune_rt_ctx *> s;
  return s;
}

struct NvtxRange {   // ranges around the time loop and the host-memref sweep
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};
}  // namespace

// ---------------------------------------------------------------------------------------------
// handles
// ---------------------------------------------------------------------------------------------
struct neptune_rt_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;       // sweeps
  cudaStream_t halo_stream = nullptr;  // halo exchange, high priority
  cudaEvent_t ev_boundary = nullptr, ev_halo = nullptr;
  cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
  double *red_partials = nullptr;   // REDUCE_BLOCKS + 1 doubles of device scratch (neptune_rt_field_reduce_sum)
  double *red_host = nullptr;       // pinned result word
  cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;  // copy engines of the pipelined host path
  std::vector<cudaEvent_t> ev_pool;                          // per-chunk events of neptune_rt_apply_host
  int64_t launches = 0;
  uint64_t next_field_id = 1;   // fields are numbered in creation order: the same numbers on every rank of an SPMD program
  std::string err;
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
  // every field and stencil created on this ctx and not yet freed: neptune_rt_finalize releases what is left, so a
  // handle owner that outlives the ctx (a neptune_b200_module) must ask neptune_rt_ctx_alive before touching them
  std::set<neptune_rt_field *> live_fields;
  std::set<neptune_rt_stencil *> live_stencils;
  // slab mode: rank-local facts are agreed once per (stencil, fields) with a 4-int ncclAllReduce(min)
  int *agree_dev = nullptr, *agree_host = nullptr;
  // slab trace (neptune_rt_slab_trace): timing events of the last traced neptune_rt_run_steps
  bool trace_on = false;
  std::vector<cudaEvent_t> trace_events;   // 4 per launch group: begin, boundary done, exchange done, interior done
  int trace_groups = 0;
  std::vector<int> trace_kind;             // per group: steps advanced * 10 + (1 = exchanged)
  std::vector<std::function<void()>> on_finalize;
  // peer-to-peer halo path (same node): ghost planes are PUSHED into the neighbour's field memory by a copy engine over
  // NVLink (cudaMemcpyAsync into CUDA-IPC-mapped memory) and ordered with stream memory operations on four flag words
  // per rank, so that no kernel — and therefore no SM — takes part in the exchange
  bool p2p = false;
  uint32_t *flags = nullptr;                       // [0] READY from up, [1] READY from down, [2] ARRIVED from up, [3] ARRIVED from down,
                                                   // [4] / [5] how many of my exported fields the up / down neighbour has unmapped (a
                                                   // watermark, see ipc_release_field), [6] / [7] that neighbour has unmapped these flags
  // Teardown of IPC-exported memory: an allocation a neighbour still maps must not be cudaFree'd. Exports towards each
  // side are numbered in the order they are made (the same numbers on both ends of the pair); the importer reports
  // "everything up to n is unmapped" into the exporter's flag words, and an allocation whose number is above that
  // waits in `deferred` (looked at again at every later free, and at finalize with a bounded wait). No collective.
  uint32_t ipc_seq[2] = {0, 0};                    // last export/import number handed out towards up / down
  std::set<uint32_t> ipc_open[2];                  // import numbers (of the up / down neighbour's fields) I still map
  struct DeferredFree { void *ptr; uint32_t need[2]; };
  std::vector<DeferredFree> deferred;
  uint32_t *peer_flags[2] = {nullptr, nullptr};    // the up / down neighbour's flag words (IPC-mapped)
  uint32_t epoch = 0;                              // exchanges so far (the same on every rank: SPMD)
  unsigned wait_flags = 0;
  CUresult (*StreamWriteValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned) = nullptr;
  CUresult (*StreamWaitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned) = nullptr;
};

struct neptune_rt_field {
  neptune_rt_ctx *ctx = nullptr;
  int dtype = NEPTUNE_RT_F64, rank = 1;
  int64_t lb[3] = {0, 0, 0}, ub[3] = {0, 0, 0};
  int64_t ext[3] = {1, 1, 1};  // canonical (z, y, x)
  void *dptr = nullptr;
  size_t bytes = 0;
  bool owned = true;   // false: caller's device memory (neptune_rt_field_wrap_device)
  // peer-to-peer halo path: the up / down neighbour's corresponding field, IPC-mapped (where my boundary planes land in
  // it depends on the ghost depth, so that is kept with the stencil's agreement, not here)
  bool peer_mapped = false;
  void *peer_base[2] = {nullptr, nullptr};
  uint32_t ipc_seq[2] = {0, 0};   // this field's export number towards the up / down neighbour (0: never exported there)
  uint64_t id = 0;     // never reused, unlike the address: what the slab agreement cache is keyed on
  size_t elsize() const { return dtype == NEPTUNE_RT_F64 ? 8 : 4; }
  int64_t numel() const { return ext[0] * ext[1] * ext[2]; }
};

struct TmapEntry {
  const void *ptr;
  int64_t nz, ny, nx;
  int box_x, box_y;
  CUtensorMap map;
};

struct neptune_rt_stencil {
  neptune_rt_ctx *ctx = nullptr;
  StencilDesc desc;
  BodyInfo info;
  const BodyEntry *body = nullptr;
  bool fma = false;
  std::string want = "auto";
  std::string want_tb = "auto";  // which temporally blocked kernel run_steps uses when tblock == 2
  int zchunk = 0;
  int tblock = 0;  // time steps fused per launch: 0 = auto (2 when the launch is eligible), 1, 2
  // slab mode (neptune_rt_stencil_set_slab): ghost planes per neighbour side of the local box and the
  // GLOBAL apply bounds along dim 0, which a temporally blocked launch needs for its level-1 planes
  bool slab_set = false;
  int64_t slab_ghost = 0, slab_glb0 = 0, slab_gub0 = 0;
  const Variant *last = nullptr;
  std::vector<TmapEntry> tmaps;
  bool slab_loop = false;   // set by neptune_rt_run_steps around its slab launches: dim-0 bounds are the GLOBAL ones
  // what the ranks agreed for the last (fields, settings) neptune_rt_run_steps saw in slab mode
  // Keyed on field ids (creation order, identical on every rank, never reused) and on every setting that feeds the
  // decision, so that all ranks hit or miss together: a miss is a collective (neighbour swaps + an all-reduce).
  // neptune_rt_stencil_set / _set_slab also drop it.
  struct Agreed {
    bool valid = false, fuse2 = false, split = false, p2p = false;
    uint64_t a = 0, b = 0;
    std::vector<uint64_t> extra;
    int tblock = 0;
    int64_t ghost = 0, glb0 = 0, gub0 = 0;
    std::string want_tb;
    // p2p: byte offsets of my boundary planes inside the up / down neighbour's copy of field a / b, for THIS ghost depth
    size_t peer_off[2][2] = {{0, 0}, {0, 0}};
  } agreed;
};

namespace {

int set_error(neptune_rt_ctx *ctx, int code, const char *fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  std::vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_last_error = buf;
  if (ctx) ctx->err = buf;
  return code;
}

#define RT_CUDA(ctx, call)                                                                                   \
  do {                                                                                                       \
    cudaError_t e_ = (call);                                                                                 \
    if (e_ != cudaSuccess)                                                                                   \
      return set_error(ctx, NEPTUNE_RT_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                       __LINE__);                                                                            \
  } while (0)

#define RT_NCCL(ctx, call)                                                                                     \
  do {                                                                                                         \
    ncclResult_t r_ = (call);                                                                                  \
    if (r_ != ncclSuccess)                                                                                     \
      return set_error(ctx, NEPTUNE_RT_ENCCL, "%s failed: %s (%s:%d)", #call, nccl().GetErrorString(r_), __FILE__, \
                       __LINE__);                                                                              \
  } while (0)

// small device helpers that are not sweeps ----------------------------------------------------
template <class T> __global__ void fill_kernel(T *p, long long n, T v) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = v;
}

// `neptune_ir.store` with bounds: subview x2 + memref.copy (DataflowLowering.cpp:181-218)
template <class T>
__global__ void box_copy_kernel(T *dst, const T *src, long long d_sz, long long d_sy, long long s_sz, long long s_sy,
                                int bz, int by, int bx) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= bx || y >= by) return;
  for (int z = blockIdx.z; z < bz; z += gridDim.z)
    dst[(long long)z * d_sz + (long long)y * d_sy + x] = src[(long long)z * s_sz + (long long)y * s_sy + x];
}

// `neptune_ir.reduce` kind = "sum" (DataflowLowering.cpp:589-698) over a sub-box.  The reference adds the
// elements one by one in row-major order; a parallel sum cannot keep that association, so this is the
// tolerance tier: fp64 accumulation for both element types, a FIXED partition (row r of the box goes
// to block r mod REDUCE_BLOCKS, element x of a row to thread x mod 256) and fixed-shape trees, so the
// result is deterministic run to run and independent of the GPU's scheduling.
constexpr int REDUCE_BLOCKS = 592;   // 4 per SM on a 148-SM B200
constexpr int REDUCE_THREADS = 256;

__device__ __forceinline__ double block_sum_256(double v, double *sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x < 32) {
    t = threadIdx.x < REDUCE_THREADS / 32 ? sh[threadIdx.x] : 0.0;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
  }
  __syncthreads();
  return t;  // valid in thread 0
}

template <class T>
__global__ void __launch_bounds__(REDUCE_THREADS) reduce_sum_rows(const T *__restrict__ src, long long sz, long long sy,
                                                                  int bz, int by, int bx, double *__restrict__ partials) {
  __shared__ double sh[REDUCE_THREADS / 32];
  double acc = 0.0;
  const long long rows = (long long)bz * by;
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const long long z = r / by, y = r - z * by;
    const T *row = src + z * sz + y * sy;
    for (int x = threadIdx.x; x < bx; x += REDUCE_THREADS) acc += (double)__ldg(row + x);
  }
  const double t = block_sum_256(acc, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

__global__ void __launch_bounds__(REDUCE_THREADS) reduce_sum_final(const double *__restrict__ partials, int n, double *__restrict__ out) {
  __shared__ double sh[REDUCE_THREADS / 32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += REDUCE_THREADS) acc += partials[i];
  const double t = block_sum_256(acc, sh);
  if (threadIdx.x == 0) *out = t;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

void canon_box(int rank, const int64_t *lb, const int64_t *ub, int64_t clb[3], int64_t cext[3]) {
  for (int a = 0; a < 3; ++a) { clb[a] = 0; cext[a] = 1; }
  for (int d = 0; d < rank; ++d) {
    clb[3 - rank + d] = lb[d];
    cext[3 - rank + d] = ub[d] - lb[d];
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------
extern "C" int neptune_rt_init(int device, neptune_rt_ctx **out) {
  if (!out) return set_error(nullptr, NEPTUNE_RT_EINVAL, "neptune_rt_init: ctx out-pointer is NULL");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0)
    return set_error(nullptr, NEPTUNE_RT_ENODEV,
                     "neptune_rt_init: no CUDA device (%s); the B200 path has no CPU fallback",
                     e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device < 0) {
    const char *lr = std::getenv("LOCAL_RANK");
    device = lr ? std::atoi(lr) % ndev : 0;
  }
  if (device >= ndev) return set_error(nullptr, NEPTUNE_RT_EINVAL, "neptune_rt_init: device %d of %d", device, ndev);
  RT_CUDA(nullptr, cudaSetDevice(device));
  neptune_rt_ctx *c = new neptune_rt_ctx();
  c->device = device;
  // any failure below releases what was built so far (neptune_rt_finalize tolerates a half-built ctx)
#define INIT_CUDA(call)                                                                                     \
  do {                                                                                                      \
    cudaError_t e_ = (call);                                                                                \
    if (e_ != cudaSuccess) {                                                                                \
      neptune_rt_finalize(c);                                                                               \
      return set_error(nullptr, NEPTUNE_RT_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e_));          \
    }                                                                                                       \
  } while (0)
  cudaDeviceProp prop;
  INIT_CUDA(cudaGetDeviceProperties(&prop, device));
  c->sm_count = prop.multiProcessorCount;
  int lo = 0, hi = 0;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  INIT_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, lo));
  INIT_CUDA(cudaStreamCreateWithPriority(&c->halo_stream, cudaStreamNonBlocking, hi));
  INIT_CUDA(cudaEventCreateWithFlags(&c->ev_boundary, cudaEventDisableTiming));
  INIT_CUDA(cudaEventCreateWithFlags(&c->ev_halo, cudaEventDisableTiming));
  INIT_CUDA(cudaEventCreate(&c->ev_t0));
  INIT_CUDA(cudaEventCreate(&c->ev_t1));
  INIT_CUDA(cudaStreamCreateWithFlags(&c->h2d_stream, cudaStreamNonBlocking));
  INIT_CUDA(cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking));
#undef INIT_CUDA
  live_ctxs().insert(c);
  *out = c;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_timer_start(neptune_rt_ctx *c) {
  if (!c) return set_error(nullptr, NEPTUNE_RT_EINVAL, "timer_start: NULL ctx");
  RT_CUDA(c, cudaEventRecord(c->ev_t0, c->stream));
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_timer_stop(neptune_rt_ctx *c, double *ms) {
  if (!c || !ms) return set_error(c, NEPTUNE_RT_EINVAL, "timer_stop: NULL argument");
  RT_CUDA(c, cudaEventRecord(c->ev_t1, c->stream));
  RT_CUDA(c, cudaEventSynchronize(c->ev_t1));
  float f = 0;
  RT_CUDA(c, cudaEventElapsedTime(&f, c->ev_t0, c->ev_t1));
  *ms = f;
  return NEPTUNE_RT_OK;
}

namespace {
// How far each neighbour has got in unmapping my exports (flag words 4 and 5, written by the neighbours).
void ipc_read_watermarks(neptune_rt_ctx *c, uint32_t wm[2]) {
  wm[0] = wm[1] = 0;
  if (c->flags && cudaMemcpy(wm, c->flags + 4, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost) != cudaSuccess) cudaGetLastError();
}
// cudaFree what no neighbour maps any more; with wait_ms > 0 poll that long for the rest, then (force) free it anyway.
void ipc_drain_deferred(neptune_rt_ctx *c, int wait_ms, bool force) {
  for (int waited = 0;; ++waited) {
    if (c->deferred.empty()) return;
    uint32_t wm[2];
    ipc_read_watermarks(c, wm);
    for (size_t i = 0; i < c->deferred.size();) {
      if (wm[0] >= c->deferred[i].need[0] && wm[1] >= c->deferred[i].need[1]) {
        cudaFree(c->deferred[i].ptr);
        c->deferred.erase(c->deferred.begin() + (long)i);
      } else {
        ++i;
      }
    }
    if (c->deferred.empty() || waited >= wait_ms) break;
    struct timespec ts = {0, 1000000};
    nanosleep(&ts, nullptr);
  }
  if (force) {
    for (auto &d : c->deferred) cudaFree(d.ptr);
    c->deferred.clear();
  }
}
// A field is going away: unmap the neighbours' copies, tell each neighbour how far my unmapping has got, and free my
// own allocation now if no neighbour maps it, later otherwise.
void ipc_release_field(neptune_rt_ctx *c, neptune_rt_field *f) {
  for (int side = 0; side < 2; ++side) {
    if (!f->peer_base[side]) continue;
    cudaIpcCloseMemHandle(f->peer_base[side]);
    f->peer_base[side] = nullptr;
    c->ipc_open[side].erase(f->ipc_seq[side]);
    // everything below the lowest number I still map is unmapped (numbers whose open failed never entered the set)
    const uint32_t wm = c->ipc_open[side].empty() ? c->ipc_seq[side] : *c->ipc_open[side].begin() - 1;
    // I am the up neighbour's DOWN neighbour (its word 5) and the down neighbour's UP neighbour (its word 4)
    if (c->peer_flags[side] && cudaMemcpy(c->peer_flags[side] + (side == 0 ? 5 : 4), &wm, sizeof wm, cudaMemcpyHostToDevice) != cudaSuccess)
      cudaGetLastError();
  }
  if (f->owned && f->dptr) {
    if (f->ipc_seq[0] || f->ipc_seq[1]) c->deferred.push_back({f->dptr, {f->ipc_seq[0], f->ipc_seq[1]}});
    else cudaFree(f->dptr);
  }
  ipc_drain_deferred(c, 0, false);
}
}  // namespace

extern "C" int neptune_rt_finalize(neptune_rt_ctx *c) {
  if (!c) return NEPTUNE_RT_OK;
  cudaSetDevice(c->device);
  for (cudaStream_t st : {c->stream, c->halo_stream, c->h2d_stream, c->d2h_stream})
    if (st) cudaStreamSynchronize(st);
  // what the caller did not free goes with the ctx (handles become invalid: neptune_rt_ctx_alive)
  while (!c->live_stencils.empty()) neptune_rt_stencil_destroy(*c->live_stencils.begin());
  while (!c->live_fields.empty()) neptune_rt_field_free(*c->live_fields.begin());
  live_ctxs().erase(c);
  if (c->comm && nccl().ok) nccl().CommDestroy(c->comm);
  for (cudaEvent_t e : {c->ev_boundary, c->ev_halo, c->ev_t0, c->ev_t1})
    if (e) cudaEventDestroy(e);
  for (cudaStream_t st : {c->stream, c->halo_stream, c->h2d_stream, c->d2h_stream})
    if (st) cudaStreamDestroy(st);
  for (cudaEvent_t e : c->ev_pool) cudaEventDestroy(e);
  for (cudaEvent_t e : c->trace_events) cudaEventDestroy(e);
  if (c->red_partials) cudaFree(c->red_partials);
  if (c->red_host) cudaFreeHost(c->red_host);
  // exported memory goes last: first what the neighbours have already unmapped, then (after at most a second of
  // waiting for ranks that are a little behind in their own teardown) the rest; the flag words themselves the same way
  ipc_drain_deferred(c, 1000, true);
  for (int side = 0; side < 2; ++side)
    if (c->peer_flags[side]) {
      const uint32_t one = 1;   // my last write through this mapping: "I have unmapped your flags" (its word 7 / 6)
      if (cudaMemcpy(c->peer_flags[side] + (side == 0 ? 7 : 6), &one, sizeof one, cudaMemcpyHostToDevice) != cudaSuccess) cudaGetLastError();
      cudaIpcCloseMemHandle(c->peer_flags[side]);
      c->peer_flags[side] = nullptr;
    }
  if (c->flags) {
    const bool expect[2] = {c->p2p && c->rank > 0, c->p2p && c->rank + 1 < c->nranks};
    for (int waited = 0; waited < 1000; ++waited) {
      uint32_t gone[2] = {0, 0};
      if (cudaMemcpy(gone, c->flags + 6, sizeof gone, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); break; }
      if ((!expect[0] || gone[0]) && (!expect[1] || gone[1])) break;
      struct timespec ts = {0, 1000000};
      nanosleep(&ts, nullptr);
    }
    cudaFree(c->flags);
  }
  if (c->agree_dev) cudaFree(c->agree_dev);
  if (c->agree_host) cudaFreeHost(c->agree_host);
  for (auto &fn : c->on_finalize) fn();
  delete c;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_ctx_alive(neptune_rt_ctx *c) { return c && live_ctxs().count(c) ? 1 : 0; }

extern "C" int neptune_rt_sync(neptune_rt_ctx *c) {
  if (!c) return set_error(nullptr, NEPTUNE_RT_EINVAL, "neptune_rt_sync: NULL ctx");
  RT_CUDA(c, cudaStreamSynchronize(c->halo_stream));
  RT_CUDA(c, cudaStreamSynchronize(c->h2d_stream));
  RT_CUDA(c, cudaStreamSynchronize(c->stream));
  RT_CUDA(c, cudaStreamSynchronize(c->d2h_stream));
  return NEPTUNE_RT_OK;
}

extern "C" const char *neptune_rt_last_error(neptune_rt_ctx *c) { return c ? c->err.c_str() : g_last_error.c_str(); }
extern "C" int64_t neptune_rt_launch_count(neptune_rt_ctx *c) { return c ? c->launches : 0; }
extern "C" void *neptune_rt_stream(neptune_rt_ctx *c) { return c ? (void *)c->stream : nullptr; }

// ---------------------------------------------------------------------------------------------
// fields
// ---------------------------------------------------------------------------------------------
extern "C" int neptune_rt_field_alloc(neptune_rt_ctx *c, int dtype, int rank, const int64_t *lb, const int64_t *ub,
                                      neptune_rt_field **out) {
  if (!c || !out || !lb || !ub) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_field_alloc: NULL argument");
  *out = nullptr;
  if (rank < 1 || rank > NEPTUNE_RT_MAX_RANK) return set_error(c, NEPTUNE_RT_EINVAL, "field rank %d (1..3)", rank);
  if (dtype != NEPTUNE_RT_F32 && dtype != NEPTUNE_RT_F64) return set_error(c, NEPTUNE_RT_EINVAL, "field dtype %d", dtype);
  neptune_rt_field *f = new neptune_rt_field();
  f->ctx = c; f->dtype = dtype; f->rank = rank;
  for (int d = 0; d < rank; ++d) {
    if (ub[d] < lb[d]) { delete f; return set_error(c, NEPTUNE_RT_EINVAL, "field bounds inverted in dim %d", d); }
    if (ub[d] - lb[d] > 0x7fffffffLL) { delete f; return set_error(c, NEPTUNE_RT_EINVAL, "field extent too large in dim %d", d); }
    f->lb[d] = lb[d]; f->ub[d] = ub[d];
  }
  int64_t clb[3];
  canon_box(rank, lb, ub, clb, f->ext);
  f->bytes = (size_t)f->numel() * f->elsize();
  cudaSetDevice(c->device);
  cudaError_t e = cudaMalloc(&f->dptr, std::max<size_t>(f->bytes, 16) + 256);  // slack keeps vector tails in-bounds
  if (e != cudaSuccess) {
    delete f;
    return set_error(c, e == cudaErrorMemoryAllocation ? NEPTUNE_RT_ENOMEM : NEPTUNE_RT_ECUDA,
                     "cudaMalloc of a %zu-byte field failed: %s", (size_t)0 + f->bytes, cudaGetErrorString(e));
  }
  f->id = c->next_field_id++;
  c->live_fields.insert(f);
  *out = f;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_field_free(neptune_rt_field *f) {
  if (!f) return NEPTUNE_RT_OK;
  f->ctx->live_fields.erase(f);
  cudaSetDevice(f->ctx->device);
  // the halo, upload and download streams may still be reading or writing this field
  for (cudaStream_t st : {f->ctx->halo_stream, f->ctx->h2d_stream, f->ctx->stream, f->ctx->d2h_stream})
    if (st) cudaStreamSynchronize(st);
  ipc_release_field(f->ctx, f);   // unmaps the neighbours' copies; cudaFree now, or once no neighbour maps this allocation
  delete f;
  return NEPTUNE_RT_OK;
}

// A field over device memory the caller owns (a solver's work vector, a torch tensor's data_ptr): the
// matrix-free `y = A(x)` hook of SURVEY.md §8f N4 — the reference applies its operators through a MatShell
// callback on host vectors (NeptunePETScRuntime.cpp:182-230); a device-resident Krylov loop calls
// neptune_rt_apply on wrapped vectors instead, with no copy.
extern "C" int neptune_rt_field_wrap_device(neptune_rt_ctx *c, int dtype, int rank, const int64_t *lb, const int64_t *ub,
                                            void *device_ptr, neptune_rt_field **out) {
  if (!c || !out || !lb || !ub || !device_ptr) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_field_wrap_device: NULL argument");
  *out = nullptr;
  if (rank < 1 || rank > NEPTUNE_RT_MAX_RANK) return set_error(c, NEPTUNE_RT_EINVAL, "field rank %d (1..3)", rank);
  if (dtype != NEPTUNE_RT_F32 && dtype != NEPTUNE_RT_F64) return set_error(c, NEPTUNE_RT_EINVAL, "field dtype %d", dtype);
  cudaPointerAttributes attr;
  cudaSetDevice(c->device);
  if (cudaPointerGetAttributes(&attr, device_ptr) != cudaSuccess || attr.type != cudaMemoryTypeDevice || attr.device != c->device) {
    cudaGetLastError();
    return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_field_wrap_device: pointer is not device memory of GPU %d", c->device);
  }
  neptune_rt_field *f = new neptune_rt_field();
  f->ctx = c; f->dtype = dtype; f->rank = rank; f->owned = false;
  for (int d = 0; d < rank; ++d) {
    if (ub[d] < lb[d] || ub[d] - lb[d] > 0x7fffffffLL) { delete f; return set_error(c, NEPTUNE_RT_EINVAL, "field bounds in dim %d", d); }
    f->lb[d] = lb[d]; f->ub[d] = ub[d];
  }
  int64_t clb[3];
  canon_box(rank, lb, ub, clb, f->ext);
  f->bytes = (size_t)f->numel() * f->elsize();
  f->dptr = device_ptr;
  f->id = c->next_field_id++;
  c->live_fields.insert(f);
  *out = f;
  return NEPTUNE_RT_OK;
}

static int plane_range(neptune_rt_field *f, int64_t p0, int64_t p1, size_t *off, size_t *len) {
  const int cd0 = 3 - f->rank;
  const int64_t n0 = f->ext[cd0];
  if (p0 < 0 || p1 > n0 || p0 > p1) return set_error(f->ctx, NEPTUNE_RT_EINVAL, "plane range [%lld, %lld) of %lld", (long long)p0, (long long)p1, (long long)n0);
  const size_t plane = n0 ? f->bytes / (size_t)n0 : 0;
  *off = (size_t)p0 * plane;
  *len = (size_t)(p1 - p0) * plane;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_field_upload_planes(neptune_rt_field *f, const void *host, int64_t p0, int64_t p1) {
  if (!f || !host) return set_error(f ? f->ctx : nullptr, NEPTUNE_RT_EINVAL, "field_upload: NULL argument");
  size_t off, len;
  if (int rc = plane_range(f, p0, p1, &off, &len)) return rc;
  if (!len) return NEPTUNE_RT_OK;
  RT_CUDA(f->ctx, cudaMemcpyAsync((char *)f->dptr + off, (const char *)host + off, len, cudaMemcpyHostToDevice, f->ctx->stream));
  return NEPTUNE_RT_OK;
}
extern "C" int neptune_rt_field_download_planes(neptune_rt_field *f, void *host, int64_t p0, int64_t p1) {
  if (!f || !host) return set_error(f ? f->ctx : nullptr, NEPTUNE_RT_EINVAL, "field_download: NULL argument");
  size_t off, len;
  if (int rc = plane_range(f, p0, p1, &off, &len)) return rc;
  if (!len) return NEPTUNE_RT_OK;
  RT_CUDA(f->ctx, cudaMemcpyAsync((char *)host + off, (const char *)f->dptr + off, len, cudaMemcpyDeviceToHost, f->ctx->stream));
  return NEPTUNE_RT_OK;
}
extern "C" int neptune_rt_field_upload(neptune_rt_field *f, const void *host) {
  if (!f) return set_error(nullptr, NEPTUNE_RT_EINVAL, "field_upload: NULL field");
  return neptune_rt_field_upload_planes(f, host, 0, f->ext[3 - f->rank]);
}
extern "C" int neptune_rt_field_download(neptune_rt_field *f, void *host) {
  if (!f) return set_error(nullptr, NEPTUNE_RT_EINVAL, "field_download: NULL field");
  return neptune_rt_field_download_planes(f, host, 0, f->ext[3 - f->rank]);
}

extern "C" int neptune_rt_field_fill(neptune_rt_field *f, double value) {
  if (!f) return set_error(nullptr, NEPTUNE_RT_EINVAL, "field_fill: NULL field");
  const long long n = f->numel();
  if (!n) return NEPTUNE_RT_OK;
  const int grid = (int)std::min<long long>((n + 255) / 256, (long long)f->ctx->sm_count * 8);
  if (f->dtype == NEPTUNE_RT_F64) fill_kernel<double><<<grid, 256, 0, f->ctx->stream>>>((double *)f->dptr, n, value);
  else fill_kernel<float><<<grid, 256, 0, f->ctx->stream>>>((float *)f->dptr, n, (float)value);
  RT_CUDA(f->ctx, cudaGetLastError());
  f->ctx->launches++;
  return NEPTUNE_RT_OK;
}

extern "C" void *neptune_rt_field_device_ptr(neptune_rt_field *f) { return f ? f->dptr : nullptr; }
extern "C" int64_t neptune_rt_field_bytes(neptune_rt_field *f) { return f ? (int64_t)f->bytes : 0; }

extern "C" int neptune_rt_field_store(neptune_rt_field *dst, neptune_rt_field *src, const int64_t *lb, const int64_t *ub) {
  if (!dst || !src) return set_error(nullptr, NEPTUNE_RT_EINVAL, "field_store: NULL field");
  neptune_rt_ctx *c = dst->ctx;
  if (dst->rank != src->rank || dst->dtype != src->dtype) return set_error(c, NEPTUNE_RT_EINVAL, "field_store: rank/dtype mismatch");
  if (!lb || !ub) {
    // whole memref.copy (DataflowLowering.cpp:176-179): shapes must agree
    for (int a = 0; a < 3; ++a)
      if (dst->ext[a] != src->ext[a]) return set_error(c, NEPTUNE_RT_EINVAL, "field_store: extents differ");
    if (dst->dptr != src->dptr && dst->bytes)
      RT_CUDA(c, cudaMemcpyAsync(dst->dptr, src->dptr, dst->bytes, cudaMemcpyDeviceToDevice, c->stream));
    return NEPTUNE_RT_OK;
  }
  int64_t blb[3], bext[3];
  canon_box(dst->rank, lb, ub, blb, bext);
  int64_t slb[3], dlb[3], tmp[3];
  canon_box(src->rank, src->lb, src->ub, slb, tmp);
  canon_box(dst->rank, dst->lb, dst->ub, dlb, tmp);
  for (int a = 0; a < 3; ++a) {
    if (bext[a] <= 0) return NEPTUNE_RT_OK;
    if (blb[a] < slb[a] || blb[a] + bext[a] > slb[a] + src->ext[a] || blb[a] < dlb[a] || blb[a] + bext[a] > dlb[a] + dst->ext[a])
      return set_error(c, NEPTUNE_RT_EINVAL, "field_store: bounds leave the source or destination box");
  }
  const long long s_sy = src->ext[2], s_sz = src->ext[1] * src->ext[2], d_sy = dst->ext[2], d_sz = dst->ext[1] * dst->ext[2];
  const long long soff = (blb[0] - slb[0]) * s_sz + (blb[1] - slb[1]) * s_sy + (blb[2] - slb[2]);
  const long long doff = (blb[0] - dlb[0]) * d_sz + (blb[1] - dlb[1]) * d_sy + (blb[2] - dlb[2]);
  dim3 block(128, 2, 1);
  dim3 grid((unsigned)((bext[2] + 127) / 128), (unsigned)((bext[1] + 1) / 2), (unsigned)std::min<int64_t>(bext[0], 1024));
  if (grid.y > 65535u) return set_error(c, NEPTUNE_RT_EINVAL, "field_store: box too tall");
  if (dst->dtype == NEPTUNE_RT_F64)
    box_copy_kernel<double><<<grid, block, 0, c->stream>>>((double *)dst->dptr + doff, (const double *)src->dptr + soff, d_sz, d_sy, s_sz, s_sy, (int)bext[0], (int)bext[1], (int)bext[2]);
  else
    box_copy_kernel<float><<<grid, block, 0, c->stream>>>((float *)dst->dptr + doff, (const float *)src->dptr + soff, d_sz, d_sy, s_sz, s_sy, (int)bext[0], (int)bext[1], (int)bext[2]);
  RT_CUDA(c, cudaGetLastError());
  c->launches++;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_field_reduce_sum(neptune_rt_field *f, const int64_t *lb, const int64_t *ub, int allreduce, double *result) {
  if (!f || !result) return set_error(f ? f->ctx : nullptr, NEPTUNE_RT_EINVAL, "field_reduce_sum: NULL argument");
  neptune_rt_ctx *c = f->ctx;
  int64_t blb[3], bext[3], flb[3], tmp[3];
  canon_box(f->rank, lb ? lb : f->lb, ub ? ub : f->ub, blb, bext);
  canon_box(f->rank, f->lb, f->ub, flb, tmp);
  bool empty = false;
  for (int a = 0; a < 3; ++a) {
    empty |= bext[a] <= 0;
    if (bext[a] > 0 && (blb[a] < flb[a] || blb[a] + bext[a] > flb[a] + f->ext[a]))
      return set_error(c, NEPTUNE_RT_EINVAL, "field_reduce_sum: bounds leave the field's box");
  }
  cudaSetDevice(c->device);
  if (!c->red_partials) {
    RT_CUDA(c, cudaMalloc(&c->red_partials, sizeof(double) * (REDUCE_BLOCKS + 1)));
    RT_CUDA(c, cudaMallocHost(&c->red_host, sizeof(double)));
  }
  double *out = c->red_partials + REDUCE_BLOCKS;
  if (empty) {
    RT_CUDA(c, cudaMemsetAsync(out, 0, sizeof(double), c->stream));
  } else {
    const long long sy = f->ext[2], sz = f->ext[1] * f->ext[2];
    const long long off = (blb[0] - flb[0]) * sz + (blb[1] - flb[1]) * sy + (blb[2] - flb[2]);
    if (f->dtype == NEPTUNE_RT_F64)
      reduce_sum_rows<double><<<REDUCE_BLOCKS, REDUCE_THREADS, 0, c->stream>>>((const double *)f->dptr + off, sz, sy, (int)bext[0], (int)bext[1], (int)bext[2], c->red_partials);
    else
      reduce_sum_rows<float><<<REDUCE_BLOCKS, REDUCE_THREADS, 0, c->stream>>>((const float *)f->dptr + off, sz, sy, (int)bext[0], (int)bext[1], (int)bext[2], c->red_partials);
    RT_CUDA(c, cudaGetLastError());
    reduce_sum_final<<<1, REDUCE_THREADS, 0, c->stream>>>(c->red_partials, REDUCE_BLOCKS, out);
    RT_CUDA(c, cudaGetLastError());
    c->launches += 2;
  }
  if (allreduce && c->nranks > 1) {
    if (!c->comm) return set_error(c, NEPTUNE_RT_ENCCL, "field_reduce_sum: no communicator (neptune_rt_comm_init)");
    RT_NCCL(c, nccl().AllReduce(out, out, 1, ncclDouble, ncclSum, c->comm, c->stream));
  }
  RT_CUDA(c, cudaMemcpyAsync(c->red_host, out, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  RT_CUDA(c, cudaStreamSynchronize(c->stream));
  *result = f->dtype == NEPTUNE_RT_F64 ? *c->red_host : (double)(float)*c->red_host;
  return NEPTUNE_RT_OK;
}

// ---------------------------------------------------------------------------------------------
// stencils
// ---------------------------------------------------------------------------------------------
extern "C" uint64_t neptune_rt_body_key(const neptune_rt_stencil_desc *d) { return d ? neptune::body_key(*d) : 0; }
extern "C" int neptune_rt_body_compiled(uint64_t key) { return neptune::find_body(key) ? 1 : 0; }
extern "C" int neptune_rt_body_jit_compile(const neptune_rt_stencil_desc *d, double *seconds, char *err, size_t err_len) {
  if (!d) return NEPTUNE_RT_EINVAL;
  std::string e;
  int rc;
  try {
    neptune::validate_desc(*d);
    rc = neptune::jit_body(*d, 0, true, e, seconds);
  } catch (const neptune::Diagnostic &dg) {
    e = dg.what();
    rc = dg.unsupported ? NEPTUNE_RT_EUNSUPPORTED : NEPTUNE_RT_EINVAL;
  }
  if (err && err_len) std::snprintf(err, err_len, "%s", e.c_str());
  return rc;
}

extern "C" int neptune_rt_stencil_create(neptune_rt_ctx *c, const neptune_rt_stencil_desc *d, neptune_rt_stencil **out) {
  if (!c || !d || !out) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_stencil_create: NULL argument");
  *out = nullptr;
  try {
    neptune::validate_desc(*d);
  } catch (const neptune::Diagnostic &e) {
    return set_error(c, e.unsupported ? NEPTUNE_RT_EUNSUPPORTED : NEPTUNE_RT_EINVAL, "%s", e.what());
  }
  const uint64_t key = neptune::body_key(*d);
  const BodyEntry *be = neptune::find_body(key);
  if (!be) {
    // not in the prebuilt catalogue: the same hand-written kernels, compiled for this body now (runtime/jit.cpp).
    // The reference lowers any body (DataflowLowering.cpp:413-423); NEPTUNE_RT_JIT=0 restores the closed catalogue.
    const char *off = std::getenv("NEPTUNE_RT_JIT");
    std::string err;
    int rc = NEPTUNE_RT_EUNSUPPORTED;
    if (off && off[0] == '0') err = "run-time compilation is switched off (NEPTUNE_RT_JIT=0)";
    else rc = neptune::jit_body(*d, c->device, false, err, nullptr);
    be = rc == NEPTUNE_RT_OK ? neptune::find_body(key) : nullptr;
    if (!be)
      return set_error(c, rc ? rc : NEPTUNE_RT_EUNSUPPORTED,
                       "apply body with structure key 0x%016llx is not in the prebuilt sm_100a kernel catalogue (%zu bodies) and could "
                       "not be compiled at run time: %s — there is no CPU fallback",
                       (unsigned long long)key, neptune::all_bodies().size(), err.c_str());
  }
  neptune_rt_stencil *s = new neptune_rt_stencil();
  s->ctx = c;
  s->desc = StencilDesc::from_c(*d);
  s->info = neptune::analyze_body(s->desc.c);
  s->body = be;
  s->fma = (d->flags & NEPTUNE_RT_FLAG_FMA) != 0;
  c->live_stencils.insert(s);
  *out = s;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_stencil_destroy(neptune_rt_stencil *s) {
  if (s && s->ctx) s->ctx->live_stencils.erase(s);
  delete s;
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_stencil_set(neptune_rt_stencil *s, const char *key, const char *value) {
  if (!s || !key || !value) return set_error(s ? s->ctx : nullptr, NEPTUNE_RT_EINVAL, "stencil_set: NULL argument");
  const std::string k = key, v = value;
  s->agreed.valid = false;   // any setting may change what the ranks have to agree on; they all set it (SPMD)
  if (k == "variant" && v.rfind("tb2", 0) == 0) {
    bool found = false;
    for (auto &var : s->body->variants) found |= var.name == v && var.kind == neptune::VARIANT_TB2;
    if (!found) return set_error(s->ctx, NEPTUNE_RT_EINVAL, "stencil_set: this body has no temporally blocked kernel '%s'", value);
    s->want_tb = v;
    s->tblock = 2;
    return NEPTUNE_RT_OK;
  }
  if (k == "variant") {
    if (v != "auto") {
      bool found = false;
      for (auto &var : s->body->variants) found |= var.name == v && var.kind != neptune::VARIANT_TB2;
      if (!found) return set_error(s->ctx, NEPTUNE_RT_EINVAL, "stencil_set: this body has no kernel variant '%s'", value);
    }
    s->want = v;
    return NEPTUNE_RT_OK;
  }
  if (k == "zchunk") { s->zchunk = std::max(0, std::atoi(value)); return NEPTUNE_RT_OK; }
  if (k == "tblock") {
    const int t = std::atoi(value);
    // 4 and 8 are depths of the communication-avoiding block in slab mode (one exchange of tblock * radius ghost planes
    // per tblock steps, set up with neptune_rt_slab_localize(halo = tblock * radius)); a launch still fuses two steps
    if (t != 0 && t != 1 && t != 2 && t != 4 && t != 8)
      return set_error(s->ctx, NEPTUNE_RT_EINVAL, "stencil_set: tblock must be 0 (auto), 1, 2, 4 or 8");
    s->tblock = t > 2 ? 2 : t;
    return NEPTUNE_RT_OK;
  }
  if (k == "fma") { s->fma = v == "1" || v == "true"; return NEPTUNE_RT_OK; }
  return set_error(s->ctx, NEPTUNE_RT_EINVAL, "stencil_set: unknown key '%s'", key);
}

extern "C" int neptune_rt_stencil_variant_count(neptune_rt_stencil *s) {
  if (!s) return 0;
  int n = 0;
  for (auto &v : s->body->variants) n += v.fma == false && v.kind != neptune::VARIANT_TB2;
  return n;
}
extern "C" const char *neptune_rt_stencil_variant_name(neptune_rt_stencil *s, int index) {
  if (!s) return nullptr;
  int n = 0;
  for (auto &v : s->body->variants)
    if (!v.fma && v.kind != neptune::VARIANT_TB2 && n++ == index) return v.name.c_str();
  return nullptr;
}

extern "C" const char *neptune_rt_stencil_variant(neptune_rt_stencil *s) {
  return (s && s->last) ? s->last->name.c_str() : "";
}

namespace {

int get_tmap(neptune_rt_stencil *s, const void *ptr, const int64_t ext[3], int box_x, int box_y, const CUtensorMap **out) {
  for (auto &e : s->tmaps)
    if (e.ptr == ptr && e.nz == ext[0] && e.ny == ext[1] && e.nx == ext[2] && e.box_x == box_x && e.box_y == box_y) {
      *out = &e.map;
      return NEPTUNE_RT_OK;
    }
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return set_error(s->ctx, NEPTUNE_RT_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  const size_t S = s->desc.c.dtype == NEPTUNE_RT_F64 ? 8 : 4;
  TmapEntry e;
  e.ptr = ptr; e.nz = ext[0]; e.ny = ext[1]; e.nx = ext[2]; e.box_x = box_x; e.box_y = box_y;
  cuuint64_t dims[3] = {(cuuint64_t)ext[2], (cuuint64_t)ext[1], (cuuint64_t)ext[0]};
  cuuint64_t strides[2] = {(cuuint64_t)ext[2] * S, (cuuint64_t)ext[2] * (cuuint64_t)ext[1] * S};
  cuuint32_t box[3] = {(cuuint32_t)box_x, (cuuint32_t)box_y, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(&e.map, S == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3,
                   const_cast<void *>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(s->ctx, NEPTUNE_RT_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  if (s->tmaps.size() >= 64) s->tmaps.erase(s->tmaps.begin());
  s->tmaps.push_back(e);
  *out = &s->tmaps.back().map;
  return NEPTUNE_RT_OK;
}

// Can the TMA-staged kernel take this launch?  It assumes every input shares the result's box
// (so that one set of coordinates addresses all of them) and 16-byte aligned rows.
bool tma_eligible(const neptune_rt_stencil *s, int n_in, neptune_rt_field *const *in, const neptune_rt_field *out) {
  const size_t S = out->elsize();
  const int VX = (int)(16 / S);
  if (out->ext[2] % VX) return false;
  if ((uintptr_t)out->dptr % 16) return false;
  for (int k = 0; k < n_in; ++k) {
    for (int a = 0; a < 3; ++a)
      if (in[k]->ext[a] != out->ext[a]) return false;
    for (int d = 0; d < out->rank; ++d)
      if (in[k]->lb[d] != out->lb[d]) return false;
    if ((uintptr_t)in[k]->dptr % 16) return false;
  }
  (void)s;
  return true;
}

// The temporally blocked kernel takes: a body with a tb2 kernel, one input sharing the result's box
// (TMA-eligible), rank 3, and apply bounds at least one point inside the box on every side (so no
// in-bounds point of either level reads outside the box; the reference's own fixtures obey this).
// In slab mode a dim-0 side that has a neighbour needs two ghost planes instead (level 1 is then
// evaluated on the first ghost plane from level-0 ghosts, exactly as the neighbour evaluates it).
bool tb2_eligible(const neptune_rt_stencil *s, neptune_rt_field *a, neptune_rt_field *b, bool slab = false,
                  int n_extra = 0, neptune_rt_field *const *extra = nullptr) {
  bool has = false;
  for (auto &v : s->body->variants) has |= v.kind == neptune::VARIANT_TB2;
  const neptune_rt_stencil_desc &d = s->desc.c;
  // inputs beyond input 0 are pointwise (the body registers a tb2 kernel only then) and stay the same
  // fields at both levels: a time loop replaces input 0 only
  if (!has || d.rank < 2 || d.n_inputs != 1 + n_extra) return false;   // rank 3: sweep_tb2 / tb2m; rank 2: sweep_rows2
  neptune_rt_field *inN[NEPTUNE_RT_MAX_INPUTS] = {a};
  for (int k = 0; k < n_extra; ++k) inN[k + 1] = extra[k];
  if (!tma_eligible(s, 1 + n_extra, inN, b) || encode_tiled() == nullptr) return false;
  for (int x = 0; x < d.rank; ++x)
    if (d.ub[x] <= d.lb[x]) return false;
  for (int x = slab ? 1 : 0; x < d.rank; ++x)
    if (d.lb[x] < d.out_lb[x] + 1 || d.ub[x] > d.out_ub[x] - 1) return false;
  if (slab) {
    const neptune_rt_ctx *c = s->ctx;
    if (!s->slab_set) return false;
    const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
    if ((up || down) && s->slab_ghost < 2) return false;
    if (!up && s->slab_glb0 < d.out_lb[0] + 1) return false;
    if (!down && s->slab_gub0 > d.out_ub[0] - 1) return false;
  }
  return true;
}

// two_steps: out = apply(apply(in)) in one temporally blocked launch (caller checked tb2_eligible).
// [g0, g1) non-empty: planes p0..g0 and g1..p1 only (equally wide: the two boundary ranges of a slab),
// in one launch where the picked kernel can take two ranges, else in two.
int do_apply(neptune_rt_stencil *s, int n_in, neptune_rt_field *const *in, neptune_rt_field *out, int64_t p0, int64_t p1,
             cudaStream_t stream, bool two_steps = false, int64_t g0 = 0, int64_t g1 = 0) {
  neptune_rt_ctx *c = s->ctx;
  const neptune_rt_stencil_desc &d = s->desc.c;
  if (n_in != d.n_inputs) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: %d inputs given, the apply has %d", n_in, d.n_inputs);
  if (!in || !out) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: NULL field");
  if (out->rank != d.rank || out->dtype != d.dtype) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: result field rank/dtype differ from the apply");
  for (int a = 0; a < d.rank; ++a)
    if (out->lb[a] != d.out_lb[a] || out->ub[a] != d.out_ub[a])
      return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: result field bounds differ from the apply's result type");
  SweepArgs a;
  for (int k = 0; k < n_in; ++k) {
    neptune_rt_field *f = in[k];
    if (!f) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: input %d is NULL", k);
    if (f->rank != d.rank || f->dtype != d.dtype) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: input %d rank/dtype differ from the apply", k);
    for (int x = 0; x < d.rank; ++x)
      if (f->lb[x] != d.in_lb[(size_t)k * d.rank + x] || f->ub[x] != d.in_ub[(size_t)k * d.rank + x])
        return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: input %d bounds differ from the apply's operand type", k);
    // byte ranges, not pointer equality: wrapped fields may be overlapping views of one work vector
    const char *ib = (const char *)f->dptr, *ob = (const char *)out->dptr;
    if (ib < ob + out->bytes && ob < ib + f->bytes)
      return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply: the result overlaps input %d in memory (the reference allocates a fresh result, DataflowLowering.cpp:273-281)", k);
  }
  const int cd0 = 3 - d.rank;
  if (p0 < 0 || p1 > out->ext[cd0] || p0 > p1) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply_planes: plane range [%lld, %lld)", (long long)p0, (long long)p1);
  if (p0 == p1 || out->numel() == 0) return NEPTUNE_RT_OK;
  const bool gap = g1 > g0;
  if (gap && (g0 <= p0 || g1 >= p1 || g0 - p0 != p1 - g1))
    return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply_planes2: [%lld, %lld) does not split [%lld, %lld) into two equally wide ranges",
                     (long long)g0, (long long)g1, (long long)p0, (long long)p1);

  const size_t S = out->elsize();
  a.out = out->dptr;
  a.out_sy = out->ext[2];
  a.out_sz = out->ext[1] * out->ext[2];
  a.nz = (int)out->ext[0]; a.ny = (int)out->ext[1]; a.nx = (int)out->ext[2];
  int64_t olb[3], tmp[3];
  canon_box(d.rank, d.out_lb, d.out_ub, olb, tmp);
  for (int k = 0; k < n_in; ++k) {
    int64_t ilb[3], iext[3];
    canon_box(d.rank, in[k]->lb, in[k]->ub, ilb, iext);
    a.in_sy[k] = iext[2];
    a.in_sz[k] = iext[1] * iext[2];
    // tap address = base + ((p + off) - in_lb) with p = q + out_lb   (DataflowLowering.cpp:381-411)
    const long long shift = (olb[0] - ilb[0]) * a.in_sz[k] + (olb[1] - ilb[1]) * a.in_sy[k] + (olb[2] - ilb[2]);
    a.in[k] = (const char *)in[k]->dptr + shift * (long long)S;
  }
  a.copy_src = in[0]->dptr;
  int64_t blb[3], bext[3];
  canon_box(d.rank, d.lb, d.ub, blb, bext);
  bool empty = false;
  for (int x = 0; x < d.rank; ++x) empty |= d.ub[x] <= d.lb[x];
  for (int x = 0; x < 3; ++x) {
    a.blo[x] = empty ? 0 : (int)(blb[x] - olb[x]);
    a.bhi[x] = empty ? 0 : (int)(blb[x] - olb[x] + bext[x]);
    a.wlo[x] = 0;
    a.whi[x] = (int)out->ext[x];
  }
  if ((two_steps || s->slab_loop) && s->slab_set) {
    // level-1 planes next to the owned slab lie on ghost planes: test them against the GLOBAL apply
    // bounds (clipped to the local box), not the owned-plane intersection the local descriptor carries
    a.blo[cd0] = (int)(std::max(s->slab_glb0, d.out_lb[0]) - d.out_lb[0]);
    a.bhi[cd0] = (int)(std::min(s->slab_gub0, d.out_ub[0]) - d.out_lb[0]);
  }
  a.wlo[cd0] = (int)p0;
  a.whi[cd0] = (int)p1;
  for (int i = 0; i < s->info.n_consts; ++i) a.consts[i] = s->info.consts[i];
  a.zchunk = s->zchunk;
  a.sm_count = c->sm_count;
  a.stream = stream;

  // pick the kernel variant
  const Variant *pick = nullptr;
  const bool tma_ok = tma_eligible(s, n_in, in, out) && encode_tiled() != nullptr;
  if (two_steps) {
    if (!tma_ok) return set_error(c, NEPTUNE_RT_EINVAL, "temporally blocked launch on fields that are not TMA-eligible");
    // auto: on wide boxes the z-marching form where the body has one (r01, 512^3: tb2m:128x14 427,
    // tb2:128x16 409, tb2:64x16 382 Gupd/s); narrow boxes keep the 64-wide tile so that fewer CTAs idle on a partial column
    if (s->want_tb == "auto") {
      const bool wide = out->ext[2] >= 256;
      // two inputs (Jacobi + rhs, r01 768^3): f64 tb2m:64x24 383 / 128x10 340 / 128x14 323 Gupd/s; f32 128x14 536 / 64x24 464
      const char *two_in = d.dtype == NEPTUNE_RT_F64 ? "tb2m:64x24" : "tb2m:128x14";
      // one input, wide box (r01, 512^3 f64): tb2p:128x14 466, tb2m:128x14 427, tb2g:128x14 415, tb2:128x16 409 Gupd/s.
      // The pipelined form pays seven planes of fill and drain per item instead of four: thin launches (the
      // boundary planes of a slab) keep tb2m.
      // r02: tb2q:128x14 (tb2p's data movement with a lean steady state, conflict-free ring pitches) 541+ Gupd/s
      const bool deep = (p1 - p0) - (g1 - g0) >= 24 && !gap;
      const char *prefer[] = {n_in > 1 ? two_in : wide ? (deep ? "tb2q:128x14" : "tb2m:128x14") : "tb2:64x16",
                              wide ? "tb2m:128x14" : "tb2:64x16"};
      for (const char *name : prefer) {
        for (auto &v : s->body->variants)
          if (v.kind == neptune::VARIANT_TB2 && v.fma == s->fma && v.name == name) { pick = &v; break; }
        if (pick) break;
      }
    }
    if (!pick)
      for (auto &v : s->body->variants)
        if (v.kind == neptune::VARIANT_TB2 && v.fma == s->fma && (s->want_tb == "auto" || v.name == s->want_tb)) { pick = &v; break; }
    if (!pick)
      for (auto &v : s->body->variants)
        if (v.kind == neptune::VARIANT_TB2 && (s->want_tb == "auto" || v.name == s->want_tb)) { pick = &v; break; }
  } else if (s->want == "auto") {
    // rank 2: the row-marching kernel (sweep_rows.cuh); narrow boxes take the narrower strip
    if (tma_ok) {
      // r01: 16384^2 9-pt rows:1024 360 / rows:512 351 / rows:256 344 Gupd/s; 4096^2 5-pt 322 / 314 / 289
      const char *name = out->ext[2] >= 2048 ? "rows:1024" : out->ext[2] >= 1024 ? "rows:512" : "rows:256";
      for (auto &v : s->body->variants)
        if (v.kind == neptune::VARIANT_ROWS && v.fma == s->fma && v.name == name) { pick = &v; break; }
      if (!pick)
        for (auto &v : s->body->variants)
          if (v.kind == neptune::VARIANT_ROWS && v.fma == s->fma) { pick = &v; break; }
    }
    if (tma_ok && !pick)
      for (auto &v : s->body->variants)
        if (v.kind == neptune::VARIANT_TMA && v.fma == s->fma) { pick = &v; break; }
    if (!pick)
      for (auto &v : s->body->variants)
        if (v.kind == neptune::VARIANT_DIRECT && v.fma == s->fma) { pick = &v; break; }
  } else {
    for (auto &v : s->body->variants)
      if (v.name == s->want && v.fma == s->fma) { pick = &v; break; }
    if (!pick)
      for (auto &v : s->body->variants)
        if (v.name == s->want) { pick = &v; break; }
    if (pick && (pick->kind == neptune::VARIANT_TMA || pick->kind == neptune::VARIANT_ROWS) && !tma_ok)
      return set_error(c, NEPTUNE_RT_EINVAL, "kernel variant '%s' needs inputs that share the result's box and 16-byte aligned rows", pick->name.c_str());
  }
  if (!pick) return set_error(c, NEPTUNE_RT_EUNSUPPORTED, "no kernel variant '%s' for this body", s->want.c_str());
  if (gap) {
    if (pick->gap_dim != cd0) {
      if (int rc = do_apply(s, n_in, in, out, p0, g0, stream, two_steps)) return rc;
      return do_apply(s, n_in, in, out, g1, p1, stream, two_steps);
    }
    a.gap_lo = (int)g0;
    a.gap_hi = (int)g1;
  }
  CUtensorMap tm_staged, tm_pw;   // copies: get_tmap's cache may move its entries on the next call
  if ((pick->kind == neptune::VARIANT_TMA || pick->kind == neptune::VARIANT_TB2) && pick->needs_tmap) {
    const CUtensorMap *tm = nullptr;
    if (int rc = get_tmap(s, in[s->body->staged]->dptr, in[s->body->staged]->ext, pick->box_x, pick->box_y, &tm)) return rc;
    tm_staged = *tm;
    a.tmap = &tm_staged;
    if (pick->kind == neptune::VARIANT_TB2 && n_in == 2) {   // the pointwise input's tiles ride in the same ring
      if (int rc = get_tmap(s, in[1]->dptr, in[1]->ext, pick->box_x, pick->box_y, &tm)) return rc;
      tm_pw = *tm;
      a.tmap_pw = &tm_pw;
    }
  }
  cudaSetDevice(c->device);
  const int e = pick->launch(a);
  if (e != 0) return set_error(c, NEPTUNE_RT_ECUDA, "launch of sweep kernel '%s' failed: %s", pick->name.c_str(), cudaGetErrorString((cudaError_t)e));
  s->last = pick;
  c->launches++;
  return NEPTUNE_RT_OK;
}

}  // namespace

extern "C" int neptune_rt_apply(neptune_rt_stencil *s, int n_in, neptune_rt_field *const *in, neptune_rt_field *out) {
  if (!s || !out) return set_error(s ? s->ctx : nullptr, NEPTUNE_RT_EINVAL, "neptune_rt_apply: NULL argument");
  return do_apply(s, n_in, in, out, 0, out->ext[3 - out->rank], s->ctx->stream);
}

extern "C" int neptune_rt_apply_planes2(neptune_rt_stencil *s, int n_in, neptune_rt_field *const *in, neptune_rt_field *out,
                                        int64_t p0, int64_t g0, int64_t g1, int64_t p1, void *stream) {
  if (!s || !out) return set_error(s ? s->ctx : nullptr, NEPTUNE_RT_EINVAL, "neptune_rt_apply_planes2: NULL argument");
  if (g1 <= g0) return set_error(s->ctx, NEPTUNE_RT_EINVAL, "neptune_rt_apply_planes2: empty gap [%lld, %lld)", (long long)g0, (long long)g1);
  return do_apply(s, n_in, in, out, p0, p1, stream ? (cudaStream_t)stream : s->ctx->stream, false, g0, g1);
}

extern "C" int neptune_rt_apply_planes(neptune_rt_stencil *s, int n_in, neptune_rt_field *const *in, neptune_rt_field *out,
                                       int64_t p0, int64_t p1, void *stream) {
  if (!s || !out) return set_error(s ? s->ctx : nullptr, NEPTUNE_RT_EINVAL, "neptune_rt_apply_planes: NULL argument");
  return do_apply(s, n_in, in, out, p0, p1, stream ? (cudaStream_t)stream : s->ctx->stream);
}

// One sweep on HOST buffers as a three-stage pipeline over dim-0 plane chunks: chunk i+1 uploads
// (copy engine, H2D) while chunk i is swept (SMs) and chunk i-1 downloads (copy engine, D2H), so a
// call costs max(upload, download) + one chunk instead of upload + sweep + download.  A chunk of
// output planes [a, b) reads input planes [a - r, b + r): its launch waits for the upload of the
// next chunk too.  The sweep of a chunk writes whole planes (copy-through outside `bounds`), so the
// downloaded result is the full result box (DataflowLowering.cpp:282-287 + :289-444).
extern "C" int neptune_rt_apply_host(neptune_rt_stencil *s, int n_in, neptune_rt_field *const *in, const void *const *in_host,
                                     neptune_rt_field *out, void *out_host, int64_t chunk_planes) {
  if (!s || !in || !in_host || !out || !out_host) return set_error(s ? s->ctx : nullptr, NEPTUNE_RT_EINVAL, "neptune_rt_apply_host: NULL argument");
  NvtxRange nvtx_range("neptune_rt_apply_host");
  neptune_rt_ctx *c = s->ctx;
  const neptune_rt_stencil_desc &d = s->desc.c;
  if (n_in != d.n_inputs) return set_error(c, NEPTUNE_RT_EINVAL, "neptune_rt_apply_host: %d inputs given, the apply has %d", n_in, d.n_inputs);
  const int cd0 = 3 - d.rank;
  const int64_t n0 = out->ext[cd0];
  int64_t radius = 0;
  for (int i = 0; i < d.n_ops; ++i)
    if (d.ops[i].opcode == NEPTUNE_RT_OP_ACCESS) radius = std::max<int64_t>(radius, std::llabs(d.ops[i].offset[0]));
  bool same_rows = true;
  for (int k = 0; k < n_in; ++k)
    same_rows &= in[k] && in[k]->rank == out->rank && in[k]->lb[0] == out->lb[0] && in[k]->ub[0] == out->ub[0];
  if (chunk_planes <= 0) {
    // about `target` chunks (pipeline fill and drain cost about 3 / target of the transfer time), at least
    // 8 MiB each so that a copy runs at full PCIe rate; NEPTUNE_RT_HOST_CHUNKS overrides the target
    static const int64_t target = [] {
      const char *e = std::getenv("NEPTUNE_RT_HOST_CHUNKS");
      const long v = e ? std::atol(e) : 0;
      return (int64_t)(v >= 3 && v <= 4096 ? v : 32);   // r01, C3: 16 -> 24.7 ms, 32 -> 23.5, 64 -> 23.6, 128 -> 25.0 (PCIe: 46 GB/s each way)
    }();
    const size_t plane = n0 ? out->bytes / (size_t)n0 : 1;
    chunk_planes = std::max<int64_t>((n0 + target - 1) / target, (int64_t)(((size_t)8 << 20) + plane - 1) / (int64_t)std::max<size_t>(plane, 1));
  }
  chunk_planes = std::max<int64_t>(chunk_planes, std::max<int64_t>(radius, 1));
  const int64_t nchunks = n0 > 0 ? (n0 + chunk_planes - 1) / chunk_planes : 0;
  if (!same_rows || nchunks < 3) {
    // not worth a pipeline (or inputs on another dim-0 box): upload, sweep, download in order
    for (int k = 0; k < n_in; ++k)
      if (in_host[k]) if (int rc = neptune_rt_field_upload(in[k], in_host[k])) return rc;
    if (int rc = neptune_rt_apply(s, n_in, in, out)) return rc;
    if (int rc = neptune_rt_field_download(out, out_host)) return rc;
    return neptune_rt_sync(c);
  }
  cudaSetDevice(c->device);
  while ((int64_t)c->ev_pool.size() < 2 * nchunks) {
    cudaEvent_t e;
    RT_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    c->ev_pool.push_back(e);
  }
  // everything queued earlier on the sweep stream (previous call's work on these fields) goes first
  RT_CUDA(c, cudaEventRecord(c->ev_boundary, c->stream));
  RT_CUDA(c, cudaStreamWaitEvent(c->h2d_stream, c->ev_boundary, 0));
  RT_CUDA(c, cudaStreamWaitEvent(c->d2h_stream, c->ev_boundary, 0));
  auto rows = [&](int64_t i, int64_t *a, int64_t *b) { *a = i * chunk_planes; *b = std::min(n0, *a + chunk_planes); };
  auto sweep = [&](int64_t i) -> int {
    int64_t a, b;
    rows(i, &a, &b);
    RT_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_pool[(size_t)std::min(i + 1, nchunks - 1)], 0));
    if (int rc = do_apply(s, n_in, in, out, a, b, c->stream)) return rc;
    RT_CUDA(c, cudaEventRecord(c->ev_pool[(size_t)(nchunks + i)], c->stream));
    RT_CUDA(c, cudaStreamWaitEvent(c->d2h_stream, c->ev_pool[(size_t)(nchunks + i)], 0));
    size_t off, len;
    if (int rc = plane_range(out, a, b, &off, &len)) return rc;
    RT_CUDA(c, cudaMemcpyAsync((char *)out_host + off, (const char *)out->dptr + off, len, cudaMemcpyDeviceToHost, c->d2h_stream));
    return NEPTUNE_RT_OK;
  };
  for (int64_t i = 0; i < nchunks; ++i) {
    int64_t a, b;
    rows(i, &a, &b);
    for (int k = 0; k < n_in; ++k) {
      if (!in_host[k]) continue;
      size_t off, len;
      if (int rc = plane_range(in[k], a, b, &off, &len)) return rc;
      RT_CUDA(c, cudaMemcpyAsync((char *)in[k]->dptr + off, (const char *)in_host[k] + off, len, cudaMemcpyHostToDevice, c->h2d_stream));
    }
    RT_CUDA(c, cudaEventRecord(c->ev_pool[(size_t)i], c->h2d_stream));
    if (i >= 1) if (int rc = sweep(i - 1)) return rc;
  }
  if (int rc = sweep(nchunks - 1)) return rc;
  return neptune_rt_sync(c);
}

// ---------------------------------------------------------------------------------------------
// slab decomposition + halos
// ---------------------------------------------------------------------------------------------
extern "C" int neptune_rt_slab_partition(int64_t n0, int nranks, int rank, int64_t *begin, int64_t *end) {
  if (nranks < 1 || rank < 0 || rank >= nranks || n0 < 0 || !begin || !end)
    return set_error(nullptr, NEPTUNE_RT_EINVAL, "slab_partition: bad arguments");
  const int64_t base = n0 / nranks, rem = n0 % nranks;
  *begin = rank * base + std::min<int64_t>(rank, rem);
  *end = *begin + base + (rank < rem ? 1 : 0);
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_slab_localize(const neptune_rt_stencil_desc *g, int nranks, int rank, int64_t halo,
                                        neptune_rt_stencil_desc *l, int64_t *lin_lb, int64_t *lin_ub) {
  if (!g || !l || !lin_lb || !lin_ub || halo < 0) return set_error(nullptr, NEPTUNE_RT_EINVAL, "slab_localize: bad arguments");
  try {
    neptune::validate_desc(*g);
  } catch (const neptune::Diagnostic &e) {
    return set_error(nullptr, e.unsupported ? NEPTUNE_RT_EUNSUPPORTED : NEPTUNE_RT_EINVAL, "%s", e.what());
  }
  int64_t b, e;
  if (int rc = neptune_rt_slab_partition(g->out_ub[0] - g->out_lb[0], nranks, rank, &b, &e)) return rc;
  const int64_t own_lo = g->out_lb[0] + b, own_hi = g->out_lb[0] + e;
  // every tap must stay inside own +- halo
  for (int i = 0; i < g->n_ops; ++i)
    if (g->ops[i].opcode == NEPTUNE_RT_OP_ACCESS && std::llabs(g->ops[i].offset[0]) > halo)
      return set_error(nullptr, NEPTUNE_RT_EINVAL, "slab_localize: halo %lld is narrower than the dim-0 stencil radius", (long long)halo);
  for (int k = 0; k < g->n_inputs; ++k)
    if (g->in_lb[(size_t)k * g->rank] != g->out_lb[0] || g->in_ub[(size_t)k * g->rank] != g->out_ub[0])
      return set_error(nullptr, NEPTUNE_RT_EUNSUPPORTED, "slab_localize: inputs must share the result's dim-0 bounds");
  *l = *g;
  std::memcpy(lin_lb, g->in_lb, sizeof(int64_t) * (size_t)g->n_inputs * g->rank);
  std::memcpy(lin_ub, g->in_ub, sizeof(int64_t) * (size_t)g->n_inputs * g->rank);
  const int64_t lo = std::max(g->out_lb[0], own_lo - halo), hi = std::min(g->out_ub[0], own_hi + halo);
  l->out_lb[0] = lo; l->out_ub[0] = hi;
  for (int k = 0; k < g->n_inputs; ++k) { lin_lb[(size_t)k * g->rank] = lo; lin_ub[(size_t)k * g->rank] = hi; }
  l->in_lb = lin_lb; l->in_ub = lin_ub;
  l->lb[0] = std::max(g->lb[0], own_lo);
  l->ub[0] = std::max(l->lb[0], std::min(g->ub[0], own_hi));
  return NEPTUNE_RT_OK;
}

namespace {
void p2p_setup(neptune_rt_ctx *c);   // below, with the rest of the peer-to-peer halo path
}

extern "C" int neptune_rt_comm_unique_id(void *id128) {
  if (!id128) return set_error(nullptr, NEPTUNE_RT_EINVAL, "comm_unique_id: NULL");
  if (!nccl().ok) return set_error(nullptr, NEPTUNE_RT_ENCCL, "NCCL (libnccl.so.2) could not be loaded");
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  RT_NCCL(nullptr, nccl().GetUniqueId(&id));
  std::memcpy(id128, &id, 128);
  return NEPTUNE_RT_OK;
}

extern "C" int neptune_rt_comm_init(neptune_rt_ctx *c, const void *id128, int nranks, int rank) {
  if (!c || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return set_error(c, NEPTUNE_RT_EINVAL, "comm_init: bad arguments");
  if (!nccl().ok) return set_error(c, NEPTUNE_RT_ENCCL, "NCCL (libnccl.so.2) could not be loaded");
  ncclUniqueId id;
  std::memcpy(&id, id128, 128);
  cudaSetDevice(c->device);
  RT_NCCL(c, nccl().CommInitRank(&c->comm, nranks, id, rank));
  c->nranks = nranks;
  c->rank = rank;
  // NCCL sets up its peer-to-peer channels on first use: do one throw-away neighbour exchange of a halo-sized
  // message now, so that the first timed exchange does not pay for connection set-up
  if (nranks > 1) {
    const size_t W = (size_t)4 << 20;
    char *buf = nullptr;
    RT_CUDA(c, cudaMalloc(&buf, 4 * W));
    RT_NCCL(c, nccl().GroupStart());
    if (rank > 0) {
      RT_NCCL(c, nccl().Send(buf, W, ncclChar, rank - 1, c->comm, c->halo_stream));
      RT_NCCL(c, nccl().Recv(buf + W, W, ncclChar, rank - 1, c->comm, c->halo_stream));
    }
    if (rank + 1 < nranks) {
      RT_NCCL(c, nccl().Send(buf + 2 * W, W, ncclChar, rank + 1, c->comm, c->halo_stream));
      RT_NCCL(c, nccl().Recv(buf + 3 * W, W, ncclChar, rank + 1, c->comm, c->halo_stream));
    }
    RT_NCCL(c, nccl().GroupEnd());
    RT_CUDA(c, cudaStreamSynchronize(c->halo_stream));
    RT_CUDA(c, cudaFree(buf));
    p2p_setup(c);
  }
  return NEPTUNE_RT_OK;
}

namespace {

// `n` bytes to and from each dim-0 neighbour (set-up traffic: IPC handles), through NCCL; synchronous
int neighbour_swap(neptune_rt_ctx *c, const void *to_up, void *from_up, const void *to_down, void *from_down, size_t n) {
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  char *buf = nullptr;
  RT_CUDA(c, cudaMalloc(&buf, 4 * n));
  if (up) RT_CUDA(c, cudaMemcpyAsync(buf, to_up, n, cudaMemcpyHostToDevice, c->halo_stream));
  if (down) RT_CUDA(c, cudaMemcpyAsync(buf + 2 * n, to_down, n, cudaMemcpyHostToDevice, c->halo_stream));
  RT_NCCL(c, nccl().GroupStart());
  if (up) {
    RT_NCCL(c, nccl().Send(buf, n, ncclChar, c->rank - 1, c->comm, c->halo_stream));
    RT_NCCL(c, nccl().Recv(buf + n, n, ncclChar, c->rank - 1, c->comm, c->halo_stream));
  }
  if (down) {
    RT_NCCL(c, nccl().Send(buf + 2 * n, n, ncclChar, c->rank + 1, c->comm, c->halo_stream));
    RT_NCCL(c, nccl().Recv(buf + 3 * n, n, ncclChar, c->rank + 1, c->comm, c->halo_stream));
  }
  RT_NCCL(c, nccl().GroupEnd());
  if (up) RT_CUDA(c, cudaMemcpyAsync(from_up, buf + n, n, cudaMemcpyDeviceToHost, c->halo_stream));
  if (down) RT_CUDA(c, cudaMemcpyAsync(from_down, buf + 3 * n, n, cudaMemcpyDeviceToHost, c->halo_stream));
  RT_CUDA(c, cudaStreamSynchronize(c->halo_stream));
  RT_CUDA(c, cudaFree(buf));
  return NEPTUNE_RT_OK;
}

// flag words + the neighbours' flag words; any failure leaves c->p2p false (the NCCL exchange is then used; the ranks
// agree on the mode in neptune_rt_run_steps)
void p2p_setup(neptune_rt_ctx *c) {
  const char *mode = std::getenv("NEPTUNE_RT_HALO");
  if (mode && std::strcmp(mode, "nccl") == 0) return;
  auto get = [](const char *name) -> void * {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
    return p;
  };
  c->StreamWriteValue32 = reinterpret_cast<decltype(c->StreamWriteValue32)>(get("cuStreamWriteValue32"));
  c->StreamWaitValue32 = reinterpret_cast<decltype(c->StreamWaitValue32)>(get("cuStreamWaitValue32"));
  if (!c->StreamWriteValue32 || !c->StreamWaitValue32) return;
  if (cudaMalloc(&c->flags, 64) != cudaSuccess) { c->flags = nullptr; cudaGetLastError(); return; }
  cudaMemset(c->flags, 0, 64);
  cudaDeviceSynchronize();
  // self-test of the stream memory operations, with the remote-write flush where the device has it
  c->wait_flags = CU_STREAM_WAIT_VALUE_GEQ | CU_STREAM_WAIT_VALUE_FLUSH;
  bool ok = c->StreamWriteValue32((CUstream)c->halo_stream, (CUdeviceptr)(c->flags + 8), 1u, CU_STREAM_WRITE_VALUE_DEFAULT) == CUDA_SUCCESS;
  if (ok && c->StreamWaitValue32((CUstream)c->halo_stream, (CUdeviceptr)(c->flags + 8), 1u, c->wait_flags) != CUDA_SUCCESS) {
    c->wait_flags = CU_STREAM_WAIT_VALUE_GEQ;
    ok = c->StreamWaitValue32((CUstream)c->halo_stream, (CUdeviceptr)(c->flags + 8), 1u, c->wait_flags) == CUDA_SUCCESS;
  }
  ok = ok && cudaStreamSynchronize(c->halo_stream) == cudaSuccess;
  cudaIpcMemHandle_t mine, from_up, from_down;
  std::memset(&from_up, 0, sizeof from_up);
  std::memset(&from_down, 0, sizeof from_down);
  ok = ok && cudaIpcGetMemHandle(&mine, c->flags) == cudaSuccess;
  if (!ok) std::memset(&mine, 0, sizeof mine);      // still take part in the swap: the neighbours are waiting for it
  char tx[128] = {0}, rx_up[128] = {0}, rx_down[128] = {0};
  tx[0] = ok ? 1 : 0;
  std::memcpy(tx + 8, &mine, sizeof mine);
  if (neighbour_swap(c, tx, rx_up, tx, rx_down, sizeof tx) != NEPTUNE_RT_OK) { cudaGetLastError(); return; }
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  if (up) {
    std::memcpy(&from_up, rx_up + 8, sizeof from_up);
    ok = ok && rx_up[0] == 1 && cudaIpcOpenMemHandle((void **)&c->peer_flags[0], from_up, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
  }
  if (down) {
    std::memcpy(&from_down, rx_down + 8, sizeof from_down);
    ok = ok && rx_down[0] == 1 && cudaIpcOpenMemHandle((void **)&c->peer_flags[1], from_down, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
  }
  cudaGetLastError();
  c->p2p = ok;
}

// the neighbours' copies of `f`, mapped into this process, and where this rank's boundary planes land in them
int p2p_map_field(neptune_rt_ctx *c, neptune_rt_field *f, int64_t own_b, int64_t own_e, int64_t ghost, bool *ok_out, size_t off_out[2]) {
  const int cd0 = 3 - f->rank;
  const size_t plane = f->ext[cd0] ? f->bytes / (size_t)f->ext[cd0] : 0;
  bool ok = f->owned;                               // wrapped memory may sit inside another allocation: not exportable as it is
  cudaIpcMemHandle_t mine;
  std::memset(&mine, 0, sizeof mine);
  ok = ok && cudaIpcGetMemHandle(&mine, f->dptr) == cudaSuccess;
  cudaGetLastError();
  // to the up neighbour: "your last owned planes go to my rows [own_b - ghost, own_b)"; to the down one: "... [own_e, own_e + ghost)"
  char tx_up[128] = {0}, tx_down[128] = {0}, rx_up[128] = {0}, rx_down[128] = {0};
  const uint64_t off_from_up = (uint64_t)(own_b - ghost) * plane, off_from_down = (uint64_t)own_e * plane;
  tx_up[0] = tx_down[0] = ok ? 1 : 0;
  std::memcpy(tx_up + 8, &mine, sizeof mine);   std::memcpy(tx_up + 8 + sizeof mine, &off_from_up, 8);
  std::memcpy(tx_down + 8, &mine, sizeof mine); std::memcpy(tx_down + 8 + sizeof mine, &off_from_down, 8);
  if (int rc = neighbour_swap(c, tx_up, rx_up, tx_down, rx_down, 128)) return rc;
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  const bool exportable = ok;   // what the neighbours were told
  for (int side = 0; side < 2; ++side) {
    if (!(side == 0 ? up : down)) continue;
    const char *rx = side == 0 ? rx_up : rx_down;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, rx + 8, sizeof h);
    uint64_t off = 0;
    std::memcpy(&off, rx + 8 + sizeof h, 8);
    // numbering: decided from the two swapped flags alone (never from a local failure), so both ends of the pair count
    // alike: the neighbour's export number for the field I am about to map is mine for the one it maps
    if (exportable && rx[0] == 1 && !f->ipc_seq[side]) f->ipc_seq[side] = ++c->ipc_seq[side];
    if (ok && rx[0] == 1 && !f->peer_base[side]) {
      if (cudaIpcOpenMemHandle(&f->peer_base[side], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { f->peer_base[side] = nullptr; ok = false; cudaGetLastError(); }
      else c->ipc_open[side].insert(f->ipc_seq[side]);
    } else if (rx[0] != 1) {
      ok = false;
    }
    off_out[side] = (size_t)off;
  }
  f->peer_mapped = ok;
  *ok_out = ok;
  return NEPTUNE_RT_OK;
}

// One exchange of `w` = ghost planes of `f` in p2p mode, in three parts:
//   announce (sweep stream, after the last launch of this rank that touches f's ghost planes): tell the neighbours that
//            they may overwrite my ghost planes;
//   push     (halo stream, after the boundary planes are computed): wait for the neighbour's announcement, copy my
//            boundary planes into its ghost planes, tell it they have arrived;
//   join     (sweep stream, before the next launch that reads ghost planes): wait for both neighbours' data.
int p2p_announce(neptune_rt_ctx *c, uint32_t e) {
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  // I am the up neighbour's DOWN neighbour: its word [1]; and the down neighbour's UP neighbour: its word [0]
  if (up && c->StreamWriteValue32((CUstream)c->stream, (CUdeviceptr)(c->peer_flags[0] + 1), e, CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS)
    return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWriteValue32 (halo announce) failed");
  if (down && c->StreamWriteValue32((CUstream)c->stream, (CUdeviceptr)(c->peer_flags[1] + 0), e, CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS)
    return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWriteValue32 (halo announce) failed");
  return NEPTUNE_RT_OK;
}
int p2p_push(neptune_rt_ctx *c, neptune_rt_field *f, const size_t peer_off[2], int64_t own_b, int64_t own_e, int64_t w, uint32_t e) {
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  const int cd0 = 3 - f->rank;
  const size_t plane = f->ext[cd0] ? f->bytes / (size_t)f->ext[cd0] : 0, n = plane * (size_t)w;
  char *base = (char *)f->dptr;
  CUstream H = (CUstream)c->halo_stream;
  if (up) {
    if (c->StreamWaitValue32(H, (CUdeviceptr)(c->flags + 0), e, c->wait_flags) != CUDA_SUCCESS) return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWaitValue32 (halo push) failed");
    RT_CUDA(c, cudaMemcpyAsync((char *)f->peer_base[0] + peer_off[0], base + (size_t)own_b * plane, n, cudaMemcpyDeviceToDevice, c->halo_stream));
    if (c->StreamWriteValue32(H, (CUdeviceptr)(c->peer_flags[0] + 3), e, CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS) return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWriteValue32 (halo arrived) failed");
  }
  if (down) {
    if (c->StreamWaitValue32(H, (CUdeviceptr)(c->flags + 1), e, c->wait_flags) != CUDA_SUCCESS) return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWaitValue32 (halo push) failed");
    RT_CUDA(c, cudaMemcpyAsync((char *)f->peer_base[1] + peer_off[1], base + (size_t)(own_e - w) * plane, n, cudaMemcpyDeviceToDevice, c->halo_stream));
    if (c->StreamWriteValue32(H, (CUdeviceptr)(c->peer_flags[1] + 2), e, CU_STREAM_WRITE_VALUE_DEFAULT) != CUDA_SUCCESS) return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWriteValue32 (halo arrived) failed");
  }
  return NEPTUNE_RT_OK;
}
int p2p_join(neptune_rt_ctx *c, uint32_t e) {
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  if (up && c->StreamWaitValue32((CUstream)c->stream, (CUdeviceptr)(c->flags + 2), e, c->wait_flags) != CUDA_SUCCESS)
    return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWaitValue32 (halo join) failed");
  if (down && c->StreamWaitValue32((CUstream)c->stream, (CUdeviceptr)(c->flags + 3), e, c->wait_flags) != CUDA_SUCCESS)
    return set_error(c, NEPTUNE_RT_ECUDA, "cuStreamWaitValue32 (halo join) failed");
  return NEPTUNE_RT_OK;
}

}  // namespace

extern "C" int neptune_rt_halo_exchange(neptune_rt_ctx *c, neptune_rt_field *f, int64_t own_begin, int64_t own_end,
                                        int64_t width, void *stream_) {
  if (!c || !f) return set_error(c, NEPTUNE_RT_EINVAL, "halo_exchange: NULL argument");
  if (c->nranks == 1 || width == 0) return NEPTUNE_RT_OK;
  if (!c->comm) return set_error(c, NEPTUNE_RT_ENCCL, "halo_exchange: no communicator (neptune_rt_comm_init)");
  cudaStream_t st = stream_ ? (cudaStream_t)stream_ : c->halo_stream;
  const int cd0 = 3 - f->rank;
  const int64_t n0 = f->ext[cd0];
  const size_t plane = n0 ? f->bytes / (size_t)n0 : 0;
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  if (own_end - own_begin < width) return set_error(c, NEPTUNE_RT_EINVAL, "halo_exchange: slab thinner than the halo");
  if ((up && own_begin < width) || (down && own_end + width > n0))
    return set_error(c, NEPTUNE_RT_EINVAL, "halo_exchange: field has no room for %lld ghost planes", (long long)width);
  char *base = (char *)f->dptr;
  const size_t n = plane * (size_t)width;
  RT_NCCL(c, nccl().GroupStart());
  if (up) {
    RT_NCCL(c, nccl().Send(base + (size_t)own_begin * plane, n, ncclChar, c->rank - 1, c->comm, st));
    RT_NCCL(c, nccl().Recv(base + (size_t)(own_begin - width) * plane, n, ncclChar, c->rank - 1, c->comm, st));
  }
  if (down) {
    RT_NCCL(c, nccl().Send(base + (size_t)(own_end - width) * plane, n, ncclChar, c->rank + 1, c->comm, st));
    RT_NCCL(c, nccl().Recv(base + (size_t)own_end * plane, n, ncclChar, c->rank + 1, c->comm, st));
  }
  RT_NCCL(c, nccl().GroupEnd());
  return NEPTUNE_RT_OK;
}

// ---------------------------------------------------------------------------------------------
// time loop
// ---------------------------------------------------------------------------------------------
namespace {
// One launch of the slab time loop: `valid` ghost planes hold the neighbours' current state going in.  A launch that
// advances T steps consumes T * radius of them; it also computes the ghost planes that stay valid afterwards (ext),
// unless the next launch would run short — then it computes the owned planes only and its result is exchanged.
neptune_rt_slab_launch slab_next(int64_t t, int64_t nsteps, int64_t radius, bool fuse2, int64_t valid) {
  neptune_rt_slab_launch L;
  const bool two = fuse2 && t + 2 <= nsteps;
  const int64_t adv = two ? 2 * radius : radius;          // ghost planes this launch consumes
  const int64_t left = nsteps - t - (two ? 2 : 1);        // steps after this launch
  const int64_t adv_next = (fuse2 && left >= 2) ? 2 * radius : radius;
  L.steps = two ? 2 : 1;
  L.exchange = (left > 0 && valid - adv < adv_next) ? 1 : 0;
  L.ext = (!L.exchange && left > 0) ? valid - adv : 0;
  return L;
}
}  // namespace

extern "C" int64_t neptune_rt_slab_schedule(int64_t nsteps, int64_t radius, int64_t ghost, int fuse2, neptune_rt_slab_launch *plan,
                                            int64_t capacity) {
  if (nsteps < 0 || radius < 0 || ghost < radius || (fuse2 && ghost < 2 * radius)) return -1;
  int64_t n = 0, valid = ghost;
  for (int64_t t = 0; t < nsteps; ++n) {
    const neptune_rt_slab_launch L = slab_next(t, nsteps, radius, fuse2 != 0, valid);
    if (plan && n < capacity) plan[n] = L;
    valid = L.exchange ? ghost : valid - L.steps * radius;
    t += L.steps;
  }
  return n;
}

extern "C" int neptune_rt_run_steps(neptune_rt_stencil *s, neptune_rt_field *fa, neptune_rt_field *fb, int n_extra,
                                    neptune_rt_field *const *extra, int64_t nsteps, neptune_rt_field **result) {
  if (!s || !fa || !fb) return set_error(s ? s->ctx : nullptr, NEPTUNE_RT_EINVAL, "run_steps: NULL argument");
  NvtxRange nvtx_range("neptune_rt_run_steps");
  neptune_rt_ctx *c = s->ctx;
  const neptune_rt_stencil_desc &d = s->desc.c;
  if (n_extra != d.n_inputs - 1) return set_error(c, NEPTUNE_RT_EINVAL, "run_steps: %d extra inputs given, the apply has %d", n_extra, d.n_inputs - 1);
  neptune_rt_field *ins[NEPTUNE_RT_MAX_INPUTS];
  for (int k = 0; k < n_extra; ++k) ins[k + 1] = extra[k];
  neptune_rt_field *cur = fa, *nxt = fb;
  const int cd0 = 3 - d.rank;
  const int64_t n0 = fa->ext[cd0];

  if (c->nranks == 1 || !c->comm) {
    int64_t t = 0;
    if (s->tblock != 1 && tb2_eligible(s, cur, nxt, false, n_extra, extra))
      for (; t + 2 <= nsteps; t += 2) {  // two steps per launch: reads cur, writes nxt
        ins[0] = cur;
        if (int rc = do_apply(s, d.n_inputs, ins, nxt, 0, n0, c->stream, true)) return rc;
        std::swap(cur, nxt);
      }
    for (; t < nsteps; ++t) {
      ins[0] = cur;
      if (int rc = do_apply(s, d.n_inputs, ins, nxt, 0, n0, c->stream)) return rc;
      std::swap(cur, nxt);
    }
    if (result) *result = cur;
    return NEPTUNE_RT_OK;
  }

  // slab mode: the local box is [own - ghost, own + ghost) clipped to the global box, as built by
  // neptune_rt_slab_localize; ghost planes exist on the sides that have a neighbour.  A launch that
  // advances T steps (T = 1, or 2 with temporal blocking) consumes T*radius valid ghost planes.  With G
  // ghost planes a rank goes G / (T*radius) launches between exchanges: launch i of such a cycle also
  // computes the ghost planes that are still valid afterwards (redundantly with the neighbour, bit for
  // bit the same arithmetic), the last launch of the cycle computes its G owned boundary planes first
  // and their exchange overlaps the interior launch.  G = T*radius is the classic one-exchange-per-launch
  // scheme; deeper ghosts trade a few redundant planes for fewer thin boundary launches and exchanges.
  int64_t radius = 0;
  for (int i = 0; i < d.n_ops; ++i)
    if (d.ops[i].opcode == NEPTUNE_RT_OP_ACCESS) radius = std::max<int64_t>(radius, std::llabs(d.ops[i].offset[0]));
  const int64_t ghost = s->slab_set ? s->slab_ghost : radius;
  // (ghost and radius are the same on every rank: this error is raised by all of them together)
  if (ghost < radius) return set_error(c, NEPTUNE_RT_EINVAL, "run_steps: %lld ghost planes are fewer than the dim-0 stencil radius", (long long)ghost);
  const bool up = c->rank > 0, down = c->rank + 1 < c->nranks;
  const int64_t own_b = up ? ghost : 0, own_e = n0 - (down ? ghost : 0);

  // Facts that can differ between ranks (slab thickness when n0 % nranks != 0, alignment of wrapped device
  // memory, ...) must not steer the exchange pattern locally: a rank that took another branch would leave
  // its neighbour blocked in ncclRecv.  They are agreed once per (fields, settings) with one ncclAllReduce(min).
  {
    neptune_rt_stencil::Agreed &ag = s->agreed;
    bool same = ag.valid && ag.tblock == s->tblock && ag.want_tb == s->want_tb && (int)ag.extra.size() == n_extra &&
                ag.ghost == ghost && ag.glb0 == s->slab_glb0 && ag.gub0 == s->slab_gub0 &&
                ((ag.a == fa->id && ag.b == fb->id) || (ag.a == fb->id && ag.b == fa->id));
    for (int k = 0; same && k < n_extra; ++k) same = ag.extra[(size_t)k] == extra[k]->id;
    if (!same) {
      int local[4];
      local[0] = (s->tblock != 1 && ghost >= 2 * radius && own_e - own_b >= 4 * radius && tb2_eligible(s, cur, nxt, true, n_extra, extra)) ? 1 : 0;
      local[1] = own_e - own_b >= 2 * ghost ? 1 : 0;    // boundary planes and interior can be split
      local[2] = own_e - own_b >= ghost ? 1 : 0;        // the owned planes cover a whole halo
      // peer-to-peer exchange: every field that has ghost planes to refresh must be mappable on every rank (the swaps
      // inside p2p_map_field are collective with the neighbours, so every rank makes them, able or not)
      cudaSetDevice(c->device);
      bool p2p_ok = c->p2p;
      {
        const char *mode = std::getenv("NEPTUNE_RT_HALO");
        if (!(mode && std::strcmp(mode, "nccl") == 0)) {
          std::vector<neptune_rt_field *> fs = {fa, fb};
          for (int k = 0; k < n_extra; ++k) fs.push_back(extra[k]);
          for (size_t i = 0; i < fs.size(); ++i) {
            bool ok = false;
            size_t off[2] = {0, 0};
            if (int rc = p2p_map_field(c, fs[i], own_b, own_e, ghost, &ok, off)) return rc;
            if (i < 2) { ag.peer_off[i][0] = off[0]; ag.peer_off[i][1] = off[1]; }   // only the state fields are pushed
            p2p_ok = p2p_ok && ok;
          }
        } else {
          p2p_ok = false;
        }
      }
      local[3] = p2p_ok ? 1 : 0;
      if (!c->agree_dev) {
        RT_CUDA(c, cudaMalloc(&c->agree_dev, 4 * sizeof(int)));
        RT_CUDA(c, cudaMallocHost(&c->agree_host, 4 * sizeof(int)));
      }
      std::memcpy(c->agree_host, local, sizeof local);
      RT_CUDA(c, cudaMemcpyAsync(c->agree_dev, c->agree_host, sizeof local, cudaMemcpyHostToDevice, c->stream));
      RT_NCCL(c, nccl().AllReduce(c->agree_dev, c->agree_dev, 4, ncclInt, ncclMin, c->comm, c->stream));
      RT_CUDA(c, cudaMemcpyAsync(c->agree_host, c->agree_dev, sizeof local, cudaMemcpyDeviceToHost, c->stream));
      RT_CUDA(c, cudaStreamSynchronize(c->stream));
      if (!c->agree_host[2])
        return set_error(c, NEPTUNE_RT_EINVAL, "run_steps: some rank's slab is thinner than the %lld ghost planes it must send (this rank owns %lld planes)",
                         (long long)ghost, (long long)(own_e - own_b));
      ag.valid = true;
      ag.fuse2 = c->agree_host[0] != 0;
      ag.split = c->agree_host[1] != 0;
      ag.p2p = c->agree_host[3] != 0;
      ag.a = fa->id; ag.b = fb->id;
      ag.extra.clear();
      for (int k = 0; k < n_extra; ++k) ag.extra.push_back(extra[k]->id);
      ag.ghost = ghost; ag.glb0 = s->slab_glb0; ag.gub0 = s->slab_gub0;
      ag.tblock = s->tblock;
      ag.want_tb = s->want_tb;
    }
  }
  const bool fuse2 = s->agreed.fuse2, split = s->agreed.split, p2p = s->agreed.p2p;

  struct LoopGuard {   // dim-0 bounds of every launch below are the global apply bounds (ghost planes are computed too)
    neptune_rt_stencil *s;
    ~LoopGuard() { s->slab_loop = false; }
  } guard{s};
  s->slab_loop = true;

  // trace: four timing events per launch group
  const bool trace = c->trace_on;
  if (trace) { c->trace_groups = 0; c->trace_kind.clear(); }
  auto trace_mark = [&](int which, cudaStream_t st) -> int {
    if (!trace) return NEPTUNE_RT_OK;
    const size_t idx = (size_t)c->trace_groups * 4 + (size_t)which;
    while (c->trace_events.size() <= idx) {
      cudaEvent_t e;
      RT_CUDA(c, cudaEventCreate(&e));
      c->trace_events.push_back(e);
    }
    RT_CUDA(c, cudaEventRecord(c->trace_events[idx], st));
    return NEPTUNE_RT_OK;
  };
  auto nccl_health = [&]() -> int {   // a dead peer must become an error, not a hung cudaStreamSynchronize
    if (!nccl().CommGetAsyncError) return NEPTUNE_RT_OK;
    ncclResult_t st = ncclSuccess;
    if (nccl().CommGetAsyncError(c->comm, &st) != ncclSuccess || (st != ncclSuccess && st != ncclInProgress))
      return set_error(c, NEPTUNE_RT_ENCCL, "NCCL reported an asynchronous error in the halo exchange: %s", nccl().GetErrorString(st));
    return NEPTUNE_RT_OK;
  };

  // One exchange of the `ghost` boundary planes of f, started after the launch that produced them (sweep stream) and
  // joined before the next launch that reads ghost planes.  p2p: copy engines + flags (no SM takes part: an NCCL
  // send/recv kernel cannot become resident beside the persistent interior launch, which holds every SM's shared memory
  // and registers — r02 trace: the NCCL exchange of a step pair completed only as the interior launch drained);
  // otherwise ncclSend/ncclRecv.
  uint32_t ex_epoch = 0;
  auto exchange_start = [&](neptune_rt_field *f) -> int {
    if (p2p) {
      ex_epoch = ++c->epoch;
      if (int rc = p2p_announce(c, ex_epoch)) return rc;
    }
    RT_CUDA(c, cudaEventRecord(c->ev_boundary, c->stream));
    RT_CUDA(c, cudaStreamWaitEvent(c->halo_stream, c->ev_boundary, 0));
    if (p2p) {
      if (int rc = p2p_push(c, f, s->agreed.peer_off[f->id == s->agreed.a ? 0 : 1], own_b, own_e, ghost, ex_epoch)) return rc;
    } else {
      if (int rc = neptune_rt_halo_exchange(c, f, own_b, own_e, ghost, c->halo_stream)) return rc;
    }
    if (int rc = trace_mark(2, c->halo_stream)) return rc;
    RT_CUDA(c, cudaEventRecord(c->ev_halo, c->halo_stream));
    return NEPTUNE_RT_OK;
  };
  auto exchange_join = [&]() -> int {
    RT_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_halo, 0));   // my own copies / sends are done: the planes may be rewritten
    if (p2p) return p2p_join(c, ex_epoch);                       // the neighbours' planes have arrived
    return NEPTUNE_RT_OK;
  };

  // every ghost plane of the first input state
  if (int rc = exchange_start(cur)) return rc;
  if (int rc = exchange_join()) return rc;
  // launches that compute ghost planes read the other inputs (coefficients, right-hand side) there:
  // bring their ghost planes up to date once (they do not change over the time loop)
  if (ghost > radius || fuse2)
    for (int k = 0; k < n_extra; ++k) {
      if (int rc = exchange_start(extra[k])) return rc;
      if (int rc = exchange_join()) return rc;
    }
  int64_t valid = ghost;   // ghost planes of `cur` that hold the neighbour's current state
  for (int64_t t = 0; t < nsteps;) {
    const neptune_rt_slab_launch L = slab_next(t, nsteps, radius, fuse2, valid);   // the plan neptune_rt_slab_schedule prints
    const bool two = L.steps == 2, exchange = L.exchange != 0;
    const int64_t adv = L.steps * radius;
    ins[0] = cur;
    if (int rc = trace_mark(0, c->stream)) return rc;
    if (!exchange) {
      // one launch over the owned planes plus the ghost planes that stay valid
      const int64_t ext = L.ext;
      if (int rc = do_apply(s, d.n_inputs, ins, nxt, own_b - (up ? ext : 0), own_e + (down ? ext : 0), c->stream, two)) return rc;
      valid -= adv;
      if (trace) { for (int w = 1; w < 4; ++w) if (int rc = trace_mark(w, c->stream)) return rc; }
    } else if (!split) {
      // thin slab (agreed by all ranks): sweep, then exchange, not overlapped
      if (int rc = do_apply(s, d.n_inputs, ins, nxt, own_b, own_e, c->stream, two)) return rc;
      if (int rc = trace_mark(1, c->stream)) return rc;
      if (int rc = exchange_start(nxt)) return rc;
      if (int rc = exchange_join()) return rc;
      if (int rc = trace_mark(3, c->stream)) return rc;
      valid = ghost;
    } else {
      // 1. the `ghost` owned planes next to each neighbour, so that their exchange can start early
      const int64_t w = ghost;
      if (up && down && own_b + w < own_e - w) {   // both sides in one launch (two plane ranges)
        if (int rc = do_apply(s, d.n_inputs, ins, nxt, own_b, own_e, c->stream, two, own_b + w, own_e - w)) return rc;
      } else {
        if (up) if (int rc = do_apply(s, d.n_inputs, ins, nxt, own_b, own_b + w, c->stream, two)) return rc;
        if (down) if (int rc = do_apply(s, d.n_inputs, ins, nxt, own_e - w, own_e, c->stream, two)) return rc;
      }
      if (int rc = trace_mark(1, c->stream)) return rc;
      // 2. halo exchange of the new state on the high-priority stream, concurrent with 3.
      if (int rc = exchange_start(nxt)) return rc;
      // 3. interior
      if (int rc = do_apply(s, d.n_inputs, ins, nxt, own_b + (up ? w : 0), own_e - (down ? w : 0), c->stream, two)) return rc;
      if (int rc = trace_mark(3, c->stream)) return rc;
      // 4. join
      if (int rc = exchange_join()) return rc;
      valid = ghost;
    }
    if (trace) { c->trace_kind.push_back((two ? 20 : 10) + (exchange ? 1 : 0)); c->trace_groups++; }
    if (exchange && !p2p) if (int rc = nccl_health()) return rc;
    std::swap(cur, nxt);
    t += two ? 2 : 1;
  }
  if (result) *result = cur;
  return NEPTUNE_RT_OK;
}

// Per-launch-group timestamps of the slab time loop (milliseconds since the first group began): begin, boundary
// planes done, exchange done, interior done.  Enable, run neptune_rt_run_steps, read.  Costs four event records per group.
extern "C" int neptune_rt_slab_trace(neptune_rt_ctx *c, int enable) {
  if (!c) return set_error(nullptr, NEPTUNE_RT_EINVAL, "slab_trace: NULL ctx");
  c->trace_on = enable != 0;
  return NEPTUNE_RT_OK;
}
extern "C" int neptune_rt_slab_trace_read(neptune_rt_ctx *c, double *out, int capacity_groups, int *kinds) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  const int n = std::min(c->trace_groups, capacity_groups);
  if (c->trace_groups == 0 || !out) return c->trace_groups;
  cudaStreamSynchronize(c->stream);
  cudaStreamSynchronize(c->halo_stream);
  for (int g = 0; g < n; ++g) {
    for (int w = 0; w < 4; ++w) {
      float ms = 0;
      cudaEventElapsedTime(&ms, c->trace_events[0], c->trace_events[(size_t)g * 4 + (size_t)w]);
      out[g * 4 + w] = ms;
    }
    if (kinds) kinds[g] = c->trace_kind[(size_t)g];
  }
  return c->trace_groups;
}

extern "C" const char *neptune_rt_stencil_halo_mode(neptune_rt_stencil *s) {
  if (!s || !s->agreed.valid) return "";
  return s->agreed.p2p ? "p2p" : "nccl";
}

extern "C" int neptune_rt_stencil_set_slab(neptune_rt_stencil *s, int64_t ghost, int64_t global_lb0, int64_t global_ub0) {
  if (!s || ghost < 0) return set_error(s ? s->ctx : nullptr, NEPTUNE_RT_EINVAL, "stencil_set_slab: bad arguments");
  s->slab_set = true;
  s->slab_ghost = ghost;
  s->slab_glb0 = global_lb0;
  s->slab_gub0 = global_ub0;
  s->agreed.valid = false;   // ghost depth and global bounds feed the agreement (every rank calls this: SPMD)
  return NEPTUNE_RT_OK;
}

// ---------------------------------------------------------------------------------------------
// memref-level entry point for MLIR-generated code
// ---------------------------------------------------------------------------------------------
static neptune_rt_ctx *default_ctx() {
  static neptune_rt_ctx *ctx = nullptr;
  if (!ctx && neptune_rt_init(-1, &ctx) != NEPTUNE_RT_OK) {
    std::fprintf(stderr, "[NeptuneRT-B200] %s\n", neptune_rt_last_error(nullptr));
    std::abort();
  }
  return ctx;
}

// What generated code calls (include/neptune_rt.h): MLIR memref descriptors in, one sweep, result in the output
// memref.  The stencil and the device fields are created on first use and cached by the descriptor's content
// (body, constants, bounds), so a host time loop that calls this once per step pays one pipelined
// upload | sweep | download (neptune_rt_apply_host) per call and no allocation.  Failures abort, as the reference's
// shims do (NeptunePETScRuntime.cpp:15-30): generated code does not check statuses.
namespace {

struct MemrefCacheEntry {
  uint64_t digest = 0;
  neptune_rt_stencil *stencil = nullptr;
  std::vector<neptune_rt_field *> in;
  neptune_rt_field *out = nullptr;
};
std::vector<MemrefCacheEntry> &memref_cache() {
  static std::vector<MemrefCacheEntry> v;
  return v;
}

uint64_t fnv(uint64_t h, const void *p, size_t n) {
  const unsigned char *b = (const unsigned char *)p;
  for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
  return h;
}

uint64_t desc_digest(const neptune_rt_stencil_desc &d) {
  uint64_t h = 1469598103934665603ull;
  h = fnv(h, &d.rank, sizeof d.rank); h = fnv(h, &d.dtype, sizeof d.dtype);
  h = fnv(h, &d.n_inputs, sizeof d.n_inputs); h = fnv(h, &d.yield_value, sizeof d.yield_value);
  h = fnv(h, &d.flags, sizeof d.flags);
  for (int i = 0; i < d.n_ops; ++i) {
    const neptune_rt_body_op &o = d.ops[i];
    h = fnv(h, &o.opcode, sizeof o.opcode); h = fnv(h, &o.a, sizeof o.a); h = fnv(h, &o.b, sizeof o.b);
    h = fnv(h, &o.input, sizeof o.input); h = fnv(h, o.offset, sizeof o.offset); h = fnv(h, &o.value, sizeof o.value);
  }
  h = fnv(h, d.lb, sizeof(int64_t) * d.rank); h = fnv(h, d.ub, sizeof(int64_t) * d.rank);
  h = fnv(h, d.out_lb, sizeof(int64_t) * d.rank); h = fnv(h, d.out_ub, sizeof(int64_t) * d.rank);
  h = fnv(h, d.in_lb, sizeof(int64_t) * (size_t)d.rank * d.n_inputs);
  h = fnv(h, d.in_ub, sizeof(int64_t) * (size_t)d.rank * d.n_inputs);
  return h;
}

[[noreturn]] void memref_die(neptune_rt_ctx *c, const char *what) {
  std::fprintf(stderr, "[NeptuneRT-B200] %s: %s\n", what, neptune_rt_last_error(c));
  std::abort();  // generated code does not check statuses (NeptunePETScRuntime.cpp:15-30)
}

// host pointers of contiguous row-major boxes -> one cached, pipelined sweep
void apply_memref_ptrs(const neptune_rt_stencil_desc *d, int n_in, const void *const *in_host, void *out_host) {
  neptune_rt_ctx *c = default_ctx();
  if (!d || !in_host || !out_host || n_in != d->n_inputs) { set_error(c, NEPTUNE_RT_EINVAL, "bad arguments"); memref_die(c, "neptune_rt_apply_memref"); }
  NvtxRange nvtx_range("neptune_rt_apply_memref");
  const uint64_t dg = desc_digest(*d);
  MemrefCacheEntry *e = nullptr;
  for (auto &x : memref_cache()) if (x.digest == dg) { e = &x; break; }
  if (!e) {
    MemrefCacheEntry n;
    n.digest = dg;
    if (neptune_rt_stencil_create(c, d, &n.stencil)) memref_die(c, "neptune_rt_stencil_create");
    n.in.assign((size_t)n_in, nullptr);
    for (int k = 0; k < n_in; ++k)
      if (neptune_rt_field_alloc(c, d->dtype, d->rank, d->in_lb + (size_t)k * d->rank, d->in_ub + (size_t)k * d->rank, &n.in[(size_t)k])) memref_die(c, "neptune_rt_field_alloc");
    if (neptune_rt_field_alloc(c, d->dtype, d->rank, d->out_lb, d->out_ub, &n.out)) memref_die(c, "neptune_rt_field_alloc");
    memref_cache().push_back(n);
    e = &memref_cache().back();
  }
  if (neptune_rt_apply_host(e->stencil, n_in, e->in.data(), in_host, e->out, out_host, 0)) memref_die(c, "neptune_rt_apply_host");
}

// MLIR memref descriptor (allocated, aligned, offset, sizes[R], strides[R]) -> host pointer of a contiguous row-major box
template <int R, class M> const void *memref_host_ptr(neptune_rt_ctx *c, const M &m, int dtype, const int64_t *lb, const int64_t *ub, const char *what) {
  int64_t expect = 1;
  for (int x = R - 1; x >= 0; --x) {
    if (m.sizes[x] != ub[x] - lb[x]) {
      set_error(c, NEPTUNE_RT_EINVAL, "%s: memref size %lld in dim %d, the apply's type says %lld", what, (long long)m.sizes[x], x, (long long)(ub[x] - lb[x]));
      memref_die(c, "neptune_rt_apply_memref");
    }
    if (m.sizes[x] > 1 && m.strides[x] != expect) {
      set_error(c, NEPTUNE_RT_EUNSUPPORTED, "%s: stride %lld in dim %d is not the identity layout (%lld); strided memrefs are not on the B200 path", what,
                (long long)m.strides[x], x, (long long)expect);
      memref_die(c, "neptune_rt_apply_memref");
    }
    expect *= m.sizes[x];
  }
  if (!m.aligned) { set_error(c, NEPTUNE_RT_EINVAL, "%s: NULL aligned pointer", what); memref_die(c, "neptune_rt_apply_memref"); }
  const size_t S = dtype == NEPTUNE_RT_F64 ? 8 : 4;
  return (const char *)m.aligned + (size_t)m.offset * S;
}

template <int R, class M> void apply_memref_rank(const neptune_rt_stencil_desc *d, int n_in, const M *in, const M *out) {
  neptune_rt_ctx *c = default_ctx();
  if (!d || !in || !out || d->rank != R || n_in != d->n_inputs || n_in > NEPTUNE_RT_MAX_INPUTS) {
    set_error(c, NEPTUNE_RT_EINVAL, "descriptor rank / input count do not match the call (rank %d, %d inputs)", R, n_in);
    memref_die(c, "neptune_rt_apply_memref");
  }
  const void *hp[NEPTUNE_RT_MAX_INPUTS];
  for (int k = 0; k < n_in; ++k)
    hp[k] = memref_host_ptr<R>(c, in[k], d->dtype, d->in_lb + (size_t)k * R, d->in_ub + (size_t)k * R, "input memref");
  void *op = const_cast<void *>(memref_host_ptr<R>(c, *out, d->dtype, d->out_lb, d->out_ub, "result memref"));
  apply_memref_ptrs(d, n_in, hp, op);
}

}  // namespace

extern "C" void neptune_rt_apply_memref_1d(const neptune_rt_stencil_desc *d, int n_in, const neptune_rt_memref_1d *in, const neptune_rt_memref_1d *out) {
  apply_memref_rank<1>(d, n_in, in, out);
}
extern "C" void neptune_rt_apply_memref_2d(const neptune_rt_stencil_desc *d, int n_in, const neptune_rt_memref_2d *in, const neptune_rt_memref_2d *out) {
  apply_memref_rank<2>(d, n_in, in, out);
}
extern "C" void neptune_rt_apply_memref_3d(const neptune_rt_stencil_desc *d, int n_in, const neptune_rt_memref_3d *in, const neptune_rt_memref_3d *out) {
  apply_memref_rank<3>(d, n_in, in, out);
}
extern "C" void neptune_rt_apply_memref(const neptune_rt_stencil_desc *d, int n_in, const void *const *in_host, void *out_host) {
  apply_memref_ptrs(d, n_in, in_host, out_host);
}
extern "C" void neptune_rt_memref_cache_clear(void) {
  for (auto &e : memref_cache()) {
    for (auto *f : e.in) neptune_rt_field_free(f);
    neptune_rt_field_free(e.out);
    neptune_rt_stencil_destroy(e.stencil);
  }
  memref_cache().clear();
}

// ---------------------------------------------------------------------------------------------
// packed descriptors + the call builder (what the MLIR pass of mlir/lib/Passes/ApplyToB200.cpp emits calls to)
// ---------------------------------------------------------------------------------------------
// A neptune_rt_stencil_desc holds pointers, which a compiler cannot put into a constant global without
// relocations; the pass therefore stores the descriptor as a flat, position-independent byte string
// (neptune_rt_desc_pack) in an llvm.mlir.global and hands that to the runtime.
namespace {
struct PackedHeader {
  uint32_t magic, version;
  int32_t rank, dtype, n_inputs, n_ops, yield_value, flags;
  int64_t lb[3], ub[3], out_lb[3], out_ub[3];
};
constexpr uint32_t PACK_MAGIC = 0x4e423230u;   // "NB20"

struct UnpackedDesc {
  neptune_rt_stencil_desc d;
  std::vector<neptune_rt_body_op> ops;
  std::vector<int64_t> in_lb, in_ub;
};

bool unpack_desc(const void *blob, int64_t bytes, UnpackedDesc &u) {
  if (!blob || bytes < (int64_t)sizeof(PackedHeader)) return false;
  PackedHeader h;
  std::memcpy(&h, blob, sizeof h);
  if (h.magic != PACK_MAGIC || h.version != 1 || h.rank < 1 || h.rank > 3 || h.n_inputs < 1 || h.n_inputs > NEPTUNE_RT_MAX_INPUTS || h.n_ops < 1 || h.n_ops > 4096) return false;
  const size_t nb = (size_t)h.n_inputs * h.rank;
  const size_t need = sizeof h + 2 * nb * sizeof(int64_t) + (size_t)h.n_ops * sizeof(neptune_rt_body_op);
  if ((size_t)bytes < need) return false;
  const char *p = (const char *)blob + sizeof h;
  u.in_lb.resize(nb); u.in_ub.resize(nb); u.ops.resize((size_t)h.n_ops);
  std::memcpy(u.in_lb.data(), p, nb * sizeof(int64_t)); p += nb * sizeof(int64_t);
  std::memcpy(u.in_ub.data(), p, nb * sizeof(int64_t)); p += nb * sizeof(int64_t);
  std::memcpy(u.ops.data(), p, (size_t)h.n_ops * sizeof(neptune_rt_body_op));
  neptune_rt_stencil_desc &d = u.d;
  std::memset(&d, 0, sizeof d);
  d.rank = h.rank; d.dtype = h.dtype; d.n_inputs = h.n_inputs; d.n_ops = h.n_ops; d.yield_value = h.yield_value; d.flags = h.flags;
  for (int k = 0; k < 3; ++k) { d.lb[k] = h.lb[k]; d.ub[k] = h.ub[k]; d.out_lb[k] = h.out_lb[k]; d.out_ub[k] = h.out_ub[k]; }
  d.ops = u.ops.data(); d.in_lb = u.in_lb.data(); d.in_ub = u.in_ub.data();
  return true;
}

struct PendingCall {
  const void *blob = nullptr;
  std::unique_ptr<UnpackedDesc> desc;
  const void *in_host[NEPTUNE_RT_MAX_INPUTS] = {};
};
std::vector<std::unique_ptr<PendingCall>> &call_cache() {   // one per descriptor global (keyed by its address)
  static std::vector<std::unique_ptr<PendingCall>> v;
  return v;
}
}  // namespace

extern "C" int64_t neptune_rt_desc_pack(const neptune_rt_stencil_desc *d, void *buf, int64_t capacity) {
  if (!d || d->rank < 1 || d->rank > 3 || d->n_inputs < 1 || d->n_ops < 1) return -1;
  const size_t nb = (size_t)d->n_inputs * d->rank;
  const size_t need = sizeof(PackedHeader) + 2 * nb * sizeof(int64_t) + (size_t)d->n_ops * sizeof(neptune_rt_body_op);
  if (!buf || capacity < (int64_t)need) return (int64_t)need;
  PackedHeader h;
  std::memset(&h, 0, sizeof h);
  h.magic = PACK_MAGIC; h.version = 1;
  h.rank = d->rank; h.dtype = d->dtype; h.n_inputs = d->n_inputs; h.n_ops = d->n_ops; h.yield_value = d->yield_value; h.flags = d->flags;
  for (int k = 0; k < d->rank; ++k) { h.lb[k] = d->lb[k]; h.ub[k] = d->ub[k]; h.out_lb[k] = d->out_lb[k]; h.out_ub[k] = d->out_ub[k]; }
  char *p = (char *)buf;
  std::memcpy(p, &h, sizeof h); p += sizeof h;
  std::memcpy(p, d->in_lb, nb * sizeof(int64_t)); p += nb * sizeof(int64_t);
  std::memcpy(p, d->in_ub, nb * sizeof(int64_t)); p += nb * sizeof(int64_t);
  for (int i = 0; i < d->n_ops; ++i) {   // field by field: padding bytes of the struct stay zero, the blob is reproducible
    neptune_rt_body_op o;
    std::memset(&o, 0, sizeof o);
    o.opcode = d->ops[i].opcode; o.a = d->ops[i].a; o.b = d->ops[i].b; o.input = d->ops[i].input;
    for (int k = 0; k < 3; ++k) o.offset[k] = d->ops[i].offset[k];
    o.value = d->ops[i].value;
    std::memcpy(p, &o, sizeof o); p += sizeof o;
  }
  return (int64_t)need;
}

extern "C" void *neptune_rt_call_begin(const void *packed_desc, int64_t bytes) {
  neptune_rt_ctx *c = default_ctx();
  for (auto &pc : call_cache())
    if (pc->blob == packed_desc) return pc.get();
  std::unique_ptr<PendingCall> pc(new PendingCall());
  pc->blob = packed_desc;
  pc->desc.reset(new UnpackedDesc());
  if (!unpack_desc(packed_desc, bytes, *pc->desc)) { set_error(c, NEPTUNE_RT_EINVAL, "malformed packed stencil descriptor"); memref_die(c, "neptune_rt_call_begin"); }
  call_cache().push_back(std::move(pc));
  return call_cache().back().get();
}

namespace {
template <int R, class M> void call_input(void *call, int32_t k, const M *m) {
  neptune_rt_ctx *c = default_ctx();
  PendingCall *pc = (PendingCall *)call;
  if (!pc || !m || k < 0 || k >= pc->desc->d.n_inputs || pc->desc->d.rank != R) { set_error(c, NEPTUNE_RT_EINVAL, "bad call-builder input %d", (int)k); memref_die(c, "neptune_rt_call_input"); }
  const neptune_rt_stencil_desc &d = pc->desc->d;
  pc->in_host[k] = memref_host_ptr<R>(c, *m, d.dtype, d.in_lb + (size_t)k * R, d.in_ub + (size_t)k * R, "input memref");
}
template <int R, class M> void call_run(void *call, const M *out) {
  neptune_rt_ctx *c = default_ctx();
  PendingCall *pc = (PendingCall *)call;
  if (!pc || !out || pc->desc->d.rank != R) { set_error(c, NEPTUNE_RT_EINVAL, "bad call-builder result"); memref_die(c, "neptune_rt_call_run"); }
  const neptune_rt_stencil_desc &d = pc->desc->d;
  for (int k = 0; k < d.n_inputs; ++k)
    if (!pc->in_host[k]) { set_error(c, NEPTUNE_RT_EINVAL, "input %d of the apply was never bound", k); memref_die(c, "neptune_rt_call_run"); }
  void *op = const_cast<void *>(memref_host_ptr<R>(c, *out, d.dtype, d.out_lb, d.out_ub, "result memref"));
  apply_memref_ptrs(&d, d.n_inputs, pc->in_host, op);
  for (int k = 0; k < d.n_inputs; ++k) pc->in_host[k] = nullptr;
}
}  // namespace

extern "C" void neptune_rt_call_input_1d(void *call, int32_t k, const neptune_rt_memref_1d *m) { call_input<1>(call, k, m); }
extern "C" void neptune_rt_call_input_2d(void *call, int32_t k, const neptune_rt_memref_2d *m) { call_input<2>(call, k, m); }
extern "C" void neptune_rt_call_input_3d(void *call, int32_t k, const neptune_rt_memref_3d *m) { call_input<3>(call, k, m); }
extern "C" void neptune_rt_call_run_1d(void *call, const neptune_rt_memref_1d *out) { call_run<1>(call, out); }
extern "C" void neptune_rt_call_run_2d(void *call, const neptune_rt_memref_2d *out) { call_run<2>(call, out); }
extern "C" void neptune_rt_call_run_3d(void *call, const neptune_rt_memref_3d *out) { call_run<3>(call, out); }
// the names func-to-llvm gives the C-interface wrappers of functions declared with `llvm.emit_c_interface`
extern "C" void *_mlir_ciface_neptune_rt_call_begin(const void *b, int64_t n) { return neptune_rt_call_begin(b, n); }
extern "C" void _mlir_ciface_neptune_rt_call_input_1d(void *c, int32_t k, const neptune_rt_memref_1d *m) { call_input<1>(c, k, m); }
extern "C" void _mlir_ciface_neptune_rt_call_input_2d(void *c, int32_t k, const neptune_rt_memref_2d *m) { call_input<2>(c, k, m); }
extern "C" void _mlir_ciface_neptune_rt_call_input_3d(void *c, int32_t k, const neptune_rt_memref_3d *m) { call_input<3>(c, k, m); }
extern "C" void _mlir_ciface_neptune_rt_call_run_1d(void *c, const neptune_rt_memref_1d *o) { call_run<1>(c, o); }
extern "C" void _mlir_ciface_neptune_rt_call_run_2d(void *c, const neptune_rt_memref_2d *o) { call_run<2>(c, o); }
extern "C" void _mlir_ciface_neptune_rt_call_run_3d(void *c, const neptune_rt_memref_3d *o) { call_run<3>(c, o); }

// ---------------------------------------------------------------------------------------------
// neptune_rt_reduce_sum_static — the box sum for callers that may not allocate (include/neptune_rt.h).
//
// neptune_rt_field_reduce_sum keeps its per-CTA partial sums in ctx scratch that it cudaMallocs on
// first use, and hands the result back through a cudaMallocHost'ed double. That is fine for a program
// that lets this library own memory. It is not fine for the callers this path is bolted on to: a
// Krylov solver (PETSc in the reference) that takes a norm or a dot product every iteration out of a
// device pool it sized itself at start-up and filled to the last byte, code emitted by a compiler
// pipeline whose every buffer comes from its own bufferisation, and a slab time loop that must not be
// stalled half-way through (cudaMalloc and cudaFree synchronise the whole device, so a first-use
// allocation inside a solver iteration also breaks the halo/interior overlap of the step around it).
// For them the partial sums live in the array below: part of the library image, so it is in place as
// soon as a GPU has loaded the module and stays until the process exits, and this entry point never
// asks CUDA (or anyone else) for memory. No ctx, no field handle: plain pointers in, a device double out.
//
// Shared state, so the rules are: one array per GPU for the whole process; calls that can overlap on
// one GPU (different streams, different host threads) race on it and must be ordered by the caller;
// nothing in it has to be zero beforehand because every call's first kernel writes all
// REDUCE_BLOCKS slots (a CTA without rows writes 0.0) before the second kernel reads them, so whatever
// the previous call left behind is overwritten, never summed.
// ---------------------------------------------------------------------------------------------
namespace {
static __device__ __align__(256) double g_reduce_static_scratch[REDUCE_BLOCKS];
static_assert(sizeof(double) * REDUCE_BLOCKS == NEPTUNE_RT_REDUCE_STATIC_SCRATCH_BYTES,
              "include/neptune_rt.h states the size of the static reduce scratch");

constexpr int STATIC_SCRATCH_MAX_DEVICES = 64;
// g_reduce_static_scratch in HOST code is only a handle for the runtime's symbol table, not an address a kernel can
// use: the device address is looked up once per GPU, with that GPU current, and kept here.
double *g_reduce_static_addr[STATIC_SCRATCH_MAX_DEVICES] = {};
std::mutex g_reduce_static_mutex;

// Same arithmetic as reduce_sum_final; the f32 rounding neptune_rt_field_reduce_sum does on the host after its
// download happens here, because this variant leaves the result on the device.
__global__ void __launch_bounds__(REDUCE_THREADS) reduce_sum_final_static(const double *__restrict__ partials, int n,
                                                                          double *__restrict__ out, int round_to_f32) {
  __shared__ double sh[REDUCE_THREADS / 32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += REDUCE_THREADS) acc += partials[i];
  const double t = block_sum_256(acc, sh);
  if (threadIdx.x == 0) *out = round_to_f32 ? (double)(float)t : t;
}

// Device address of the scratch array on the CURRENT device (first use per device: symbol lookup, which also makes
// the runtime load this module on that device if it has not yet).
int reduce_static_scratch(int device, double **out) {
  std::lock_guard<std::mutex> lock(g_reduce_static_mutex);
  if (!g_reduce_static_addr[device]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_reduce_static_scratch);
    if (e != cudaSuccess || !p)
      return set_error(nullptr, NEPTUNE_RT_ECUDA, "reduce_sum_static: cudaGetSymbolAddress of the scratch array failed on device %d: %s",
                       device, cudaGetErrorString(e));
    g_reduce_static_addr[device] = (double *)p;
  }
  *out = g_reduce_static_addr[device];
  return NEPTUNE_RT_OK;
}

// A pointer the kernels may dereference on `device`: device or managed memory of that GPU.
bool usable_on_device(const void *p, int device) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  if (a.type == cudaMemoryTypeManaged) return true;
  return a.type == cudaMemoryTypeDevice && a.device == device;
}
}  // namespace

extern "C" int neptune_rt_reduce_sum_static(const void *src, int dtype, int rank, const int64_t *shape, const int64_t *lb,
                                            const int64_t *ub, double *result_dev, void *stream) {
  if (!src || !shape || !result_dev) return set_error(nullptr, NEPTUNE_RT_EINVAL, "reduce_sum_static: NULL src, shape or result_dev");
  if (dtype != NEPTUNE_RT_F32 && dtype != NEPTUNE_RT_F64)
    return set_error(nullptr, NEPTUNE_RT_EINVAL, "reduce_sum_static: dtype %d is neither NEPTUNE_RT_F32 nor NEPTUNE_RT_F64", dtype);
  if (rank < 1 || rank > NEPTUNE_RT_MAX_RANK)
    return set_error(nullptr, NEPTUNE_RT_EINVAL, "reduce_sum_static: rank %d outside 1..%d", rank, NEPTUNE_RT_MAX_RANK);
  int64_t zero[3] = {0, 0, 0}, flb[3], fext[3], blb[3], bext[3];
  for (int d = 0; d < rank; ++d)
    if (shape[d] < 0 || shape[d] > NEPTUNE_RT_REDUCE_STATIC_MAX_EXTENT)
      return set_error(nullptr, NEPTUNE_RT_EINVAL,
                       "reduce_sum_static: shape[%d] = %lld is beyond the limit of %lld elements per axis "
                       "(NEPTUNE_RT_REDUCE_STATIC_MAX_EXTENT); split the axis or use neptune_rt_field_reduce_sum",
                       d, (long long)shape[d], (long long)NEPTUNE_RT_REDUCE_STATIC_MAX_EXTENT);
  canon_box(rank, zero, shape, flb, fext);
  canon_box(rank, lb ? lb : zero, ub ? ub : shape, blb, bext);
  bool empty = false;
  for (int a = 0; a < 3; ++a) {
    empty |= bext[a] <= 0;
    if (bext[a] > 0 && (blb[a] < 0 || blb[a] + bext[a] > fext[a]))
      return set_error(nullptr, NEPTUNE_RT_EINVAL, "reduce_sum_static: [lb, ub) leaves the array's shape");
  }
  int device = -1;
  RT_CUDA(nullptr, cudaGetDevice(&device));
  if (device < 0 || device >= STATIC_SCRATCH_MAX_DEVICES)
    return set_error(nullptr, NEPTUNE_RT_EUNSUPPORTED, "reduce_sum_static: device ordinal %d outside 0..%d", device, STATIC_SCRATCH_MAX_DEVICES - 1);
  if (!usable_on_device(src, device) || !usable_on_device(result_dev, device))
    return set_error(nullptr, NEPTUNE_RT_EINVAL,
                     "reduce_sum_static: src and result_dev must be device memory of the current device (%d); "
                     "make the GPU that owns them current before the call", device);
  if (((uintptr_t)result_dev & 7) != 0) return set_error(nullptr, NEPTUNE_RT_EINVAL, "reduce_sum_static: result_dev is not 8-byte aligned");
  double *scratch = nullptr;
  if (int rc = reduce_static_scratch(device, &scratch)) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (empty) {
    RT_CUDA(nullptr, cudaMemsetAsync(result_dev, 0, sizeof(double), s));   // +0.0, as the other entry point returns
    return NEPTUNE_RT_OK;
  }
  const long long sy = fext[2], sz = fext[1] * fext[2];
  const long long off = blb[0] * sz + blb[1] * sy + blb[2];
  if (dtype == NEPTUNE_RT_F64)
    reduce_sum_rows<double><<<REDUCE_BLOCKS, REDUCE_THREADS, 0, s>>>((const double *)src + off, sz, sy, (int)bext[0], (int)bext[1], (int)bext[2], scratch);
  else
    reduce_sum_rows<float><<<REDUCE_BLOCKS, REDUCE_THREADS, 0, s>>>((const float *)src + off, sz, sy, (int)bext[0], (int)bext[1], (int)bext[2], scratch);
  RT_CUDA(nullptr, cudaGetLastError());
  reduce_sum_final_static<<<1, REDUCE_THREADS, 0, s>>>(scratch, REDUCE_BLOCKS, result_dev, dtype == NEPTUNE_RT_F32);
  RT_CUDA(nullptr, cudaGetLastError());
  return NEPTUNE_RT_OK;
}
