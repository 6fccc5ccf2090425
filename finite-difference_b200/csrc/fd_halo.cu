// Halo exchange between the slabs of neighbouring GPUs over peer memory
// (NVLink / NVSwitch), without a communication library on the data path.
//
// The reference has no counterpart (it is a single-process numpy program); the
// operation replaces the isend/irecv pair of finite_difference/_dist.py.
//
// Every rank exports its level buffers and a few flag words with CUDA IPC
// (fd_ipc_export / fd_ipc_open).  After the boundary planes of a new level
// have been computed a rank
//   1. pushes them straight into the ghost planes of its neighbours
//      (fd_halo_push: 128-bit stores into the peer's memory),
//   2. publishes the level number in the neighbours' flag words (same call:
//      a system-scope release store, stream-ordered after the copy), and
//   3. waits until its own flag words show that the neighbours' planes have
//      arrived (fd_halo_wait: one thread polling with acquire loads).
// A level number only grows, so no flag is ever reset.

#include <cuda_runtime.h>

#include <cstring>

#include "fd_common.cuh"

namespace fdb {

// copy `n16` 16-byte words per side; either side may be absent (null pointers)
__global__ void __launch_bounds__(256)
halo_push_kernel(const uint4* __restrict__ src_lo, uint4* __restrict__ dst_lo, const uint4* __restrict__ src_hi,
                 uint4* __restrict__ dst_hi, int64_t n16) {
  const uint4* src = blockIdx.y == 0 ? src_lo : src_hi;
  uint4* dst = blockIdx.y == 0 ? dst_lo : dst_hi;
  if (!src || !dst) return;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = __ldg(src + i);
}

// runs after halo_push_kernel on the same stream: everything that kernel wrote
// is complete; the fence + release store make it visible before the flag is
__global__ void halo_signal_kernel(long long* flag_a, long long* flag_b, long long value) {
  __threadfence_system();
  if (flag_a) asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(flag_a), "l"(value) : "memory");
  if (flag_b) asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(flag_b), "l"(value) : "memory");
}

// one thread polls the rank's own flag words until both have reached `value`;
// gives up after `timeout_ns` and raises FD_FLAG_HALO_TIMEOUT instead of hanging
__global__ void halo_wait_kernel(const long long* flag_a, const long long* flag_b, long long value, long long timeout_ns,
                                 long long* flags) {
  long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    long long a = value, b = value;
    if (flag_a) asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(a) : "l"(flag_a) : "memory");
    if (flag_b) asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(b) : "l"(flag_b) : "memory");
    if (a >= value && b >= value) return;
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) {
      if (flags) raise_flag(flags, FD_FLAG_HALO_TIMEOUT);
      return;
    }
    __nanosleep(200);
  }
}


// ---------------------------------------------------------------------------
// One-launch exchange: copy + publish + wait.
//
// Up to FD_MAX_LINKS fields at once.  Every block copies its share of one side
// of one field into the neighbour's ghost planes, makes its stores visible
// system-wide and takes a ticket; the block that draws the last ticket knows
// that every plane has been written: it publishes the level number in the
// neighbours' flag words and then waits for this rank's own flag words (the
// neighbours' planes arriving here).  The wait cannot dead-lock: a neighbour's
// publication depends on nothing this rank does in this launch.  The ticket
// counter is a device word the caller provides zeroed once; the last block
// resets it for the next launch on the stream.
struct HaloLinks {
  fd_halo_link_t link[FD_MAX_LINKS];
  int32_t nlinks;
  long long timeout_ns;
  long long* flags;
  unsigned* counter;
};

__global__ void __launch_bounds__(256)
halo_exchange_kernel(const __grid_constant__ HaloLinks L) {
  const fd_halo_link_t& k = L.link[blockIdx.y >> 1];
  const bool hi = blockIdx.y & 1;
  const uint4* src = (const uint4*)(hi ? k.src_hi : k.src_lo);
  uint4* dst = (uint4*)(hi ? k.peer_dst_hi : k.peer_dst_lo);
  if (src && dst) {
    const int64_t n16 = k.nbytes / 16;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x)
      dst[i] = __ldg(src + i);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x != 0) return;
  const unsigned total = gridDim.x * gridDim.y;
  if (atomicAdd(L.counter, 1u) != total - 1) return;
  *L.counter = 0u;
  __threadfence_system();
  for (int j = 0; j < L.nlinks; ++j) {
    const fd_halo_link_t& q = L.link[j];
    if (q.peer_flag_lo) asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(q.peer_flag_lo), "l"(q.value) : "memory");
    if (q.peer_flag_hi) asm volatile("st.release.sys.global.s64 [%0], %1;" ::"l"(q.peer_flag_hi), "l"(q.value) : "memory");
  }
  long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (int j = 0; j < L.nlinks; ++j) {
    const fd_halo_link_t& q = L.link[j];
    for (;;) {
      long long a = q.value, b = q.value;
      if (q.my_flag_lo) asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(a) : "l"(q.my_flag_lo) : "memory");
      if (q.my_flag_hi) asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(b) : "l"(q.my_flag_hi) : "memory");
      if (a >= q.value && b >= q.value) break;
      long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t - t0 > L.timeout_ns) {
        if (L.flags) raise_flag(L.flags, FD_FLAG_HALO_TIMEOUT);
        return;
      }
      __nanosleep(200);
    }
  }
}

static int launch_halo_exchange(const fd_halo_link_t* links, int nlinks, void* counter, long long timeout_ns, void* flags,
                                cudaStream_t st) {
  if (nlinks < 1 || nlinks > FD_MAX_LINKS) return fail("fd_halo_exchange: %d links, supported 1..%d", nlinks, FD_MAX_LINKS);
  if (!links || !counter) return fail("fd_halo_exchange: NULL argument");
  HaloLinks L;
  memset(&L, 0, sizeof L);
  long long most = 0;
  auto aligned = [](const void* p) { return ((uintptr_t)p & 15u) == 0; };
  for (int j = 0; j < nlinks; ++j) {
    const fd_halo_link_t& q = links[j];
    if (q.nbytes < 0 || q.nbytes % 16) return fail("fd_halo_exchange: %lld bytes, must be a multiple of 16", q.nbytes);
    if (!aligned(q.src_lo) || !aligned(q.peer_dst_lo) || !aligned(q.src_hi) || !aligned(q.peer_dst_hi))
      return fail("fd_halo_exchange: planes must be 16-byte aligned");
    if ((q.src_lo == nullptr) != (q.peer_dst_lo == nullptr) || (q.src_hi == nullptr) != (q.peer_dst_hi == nullptr))
      return fail("fd_halo_exchange: a side needs both its source and its destination");
    L.link[j] = q;
    if (q.nbytes > most) most = q.nbytes;
  }
  L.nlinks = nlinks;
  L.timeout_ns = timeout_ns;
  L.flags = (long long*)flags;
  L.counter = (unsigned*)counter;
  long long blocks = (most / 16 + 255) / 256;
  if (blocks > 148) blocks = 148;
  if (blocks < 1) blocks = 1;
  halo_exchange_kernel<<<dim3((unsigned)blocks, 2 * nlinks), 256, 0, st>>>(L);
  return check_launch("fd_halo_exchange");
}

// ---------------------------------------------------------------------------
// One statement of a slab step, issued natively: interior on the compute stream,
// boundary planes and their exchange on the communication stream, the events
// that order them -- one call instead of a dozen calls from Python.
struct Seq {
  cudaEvent_t start, rims, halo, side_start, side_done;
  unsigned* counter;
  bool halo_recorded;
};

template <typename T>
static int launch_piece(const fd_slab_step_t* s, int lo, int hi, void* stream);

template <>
int launch_piece<double>(const fd_slab_step_t* s, int lo, int hi, void* stream) {
  if (s->kind == 1) {
    fd_pair_t d = *(const fd_pair_t*)s->desc;
    d.box_lo[s->dim] = lo;
    d.box_hi[s->dim] = hi;
    return fd_fused_pair_f64(&d, stream);
  }
  if (s->kind == 2) {
    fd_expr_t d = *(const fd_expr_t*)s->desc;
    d.base.box_lo[s->dim] = lo;
    d.base.box_hi[s->dim] = hi;
    return fd_fused_expr_f64(&d, stream);
  }
  fd_fused_t d = *(const fd_fused_t*)s->desc;
  d.box_lo[s->dim] = lo;
  d.box_hi[s->dim] = hi;
  return fd_fused_step_f64(&d, stream);
}

template <>
int launch_piece<float>(const fd_slab_step_t* s, int lo, int hi, void* stream) {
  if (s->kind == 1) {
    fd_pair_t d = *(const fd_pair_t*)s->desc;
    d.box_lo[s->dim] = lo;
    d.box_hi[s->dim] = hi;
    return fd_fused_pair_f32(&d, stream);
  }
  if (s->kind == 2) {
    fd_expr_t d = *(const fd_expr_t*)s->desc;
    d.base.box_lo[s->dim] = lo;
    d.base.box_hi[s->dim] = hi;
    return fd_fused_expr_f32(&d, stream);
  }
  fd_fused_t d = *(const fd_fused_t*)s->desc;
  d.box_lo[s->dim] = lo;
  d.box_hi[s->dim] = hi;
  return fd_fused_step_f32(&d, stream);
}

#define FD_CUDA(call, what)                                                      \
  do {                                                                           \
    cudaError_t e_ = (call);                                                     \
    if (e_ != cudaSuccess) return fail("%s: %s", what, cudaGetErrorString(e_)); \
  } while (0)

template <typename T>
static int slab_step(void* seq, const fd_slab_step_t* s) {
  if (!seq || !s || !s->desc) return fail("fd_slab_step: NULL argument");
  if (s->kind < 0 || s->kind > 2 || s->dim < 0 || s->dim > 2) return fail("fd_slab_step: bad kind or dimension");
  const int H = s->depth;
  if (H < 1 || s->hi - s->lo < 3 * H) return fail("fd_slab_step: a slab of %d planes is too thin for %d ghost planes", s->hi - s->lo, H);
  if (s->nlinks < 0 || s->nlinks > FD_MAX_LINKS) return fail("fd_slab_step: %d links, supported 0..%d", s->nlinks, FD_MAX_LINKS);
  Seq* q = (Seq*)seq;
  cudaStream_t compute = (cudaStream_t)s->compute_stream, comm = (cudaStream_t)s->comm_stream;
  FD_CUDA(cudaEventRecord(q->start, compute), "fd_slab_step");
  // interior first: on an idle GPU it starts at once; the boundary planes go to the
  // (high-priority) communication stream and take the next free SMs
  if (launch_piece<T>(s, s->lo + H, s->hi - H, compute)) return 1;
  FD_CUDA(cudaStreamWaitEvent(comm, q->start, 0), "fd_slab_step");
  if (launch_piece<T>(s, s->lo, s->lo + H, comm)) return 1;
  if (launch_piece<T>(s, s->hi - H, s->hi, comm)) return 1;
  FD_CUDA(cudaEventRecord(q->rims, comm), "fd_slab_step");
  if (s->nlinks && launch_halo_exchange(s->link, s->nlinks, q->counter, s->timeout_ns, s->flags, comm)) return 1;
  FD_CUDA(cudaEventRecord(q->halo, comm), "fd_slab_step");
  q->halo_recorded = true;
  FD_CUDA(cudaStreamWaitEvent(compute, q->rims, 0), "fd_slab_step");   // later work on the compute stream sees the whole level
  return 0;
}

}  // namespace fdb

extern "C" {

int fd_ipc_export(const void* dev_ptr, void* handle_out, long long* offset_out) {
  using namespace fdb;
  if (!dev_ptr || !handle_out || !offset_out) return fail("fd_ipc_export: NULL argument");
  // the handle describes the whole allocation the pointer lies in: find its base
  typedef int (*get_range_t)(unsigned long long*, size_t*, unsigned long long);
  static get_range_t get_range = nullptr;
  if (!get_range) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    cudaError_t e = cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qr);
    if (e != cudaSuccess || !fn) return fail("fd_ipc_export: cuMemGetAddressRange is not available (%s)", cudaGetErrorString(e));
    get_range = (get_range_t)fn;
  }
  unsigned long long base = 0;
  size_t size = 0;
  if (get_range(&base, &size, (unsigned long long)(uintptr_t)dev_ptr) != 0) return fail("fd_ipc_export: not a device allocation");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail("fd_ipc_export: cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == FD_IPC_HANDLE_BYTES, "handle size");
  memcpy(handle_out, &h, sizeof h);
  *offset_out = (long long)((unsigned long long)(uintptr_t)dev_ptr - base);
  return 0;
}

int fd_ipc_open(const void* handle, void** base_out) {
  using namespace fdb;
  if (!handle || !base_out) return fail("fd_ipc_open: NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  cudaError_t e = cudaIpcOpenMemHandle(base_out, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail("fd_ipc_open: cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
  }
  return 0;
}

int fd_ipc_close(void* base) {
  using namespace fdb;
  cudaError_t e = cudaIpcCloseMemHandle(base);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail("fd_ipc_close: %s", cudaGetErrorString(e));
  }
  return 0;
}

int fd_halo_push(const void* src_lo, void* peer_dst_lo, const void* src_hi, void* peer_dst_hi, long long nbytes,
                 void* peer_flag_lo, void* peer_flag_hi, long long value, void* stream) {
  using namespace fdb;
  if (nbytes < 0 || nbytes % 16) return fail("fd_halo_push: %lld bytes, must be a multiple of 16", nbytes);
  auto aligned = [](const void* p) { return ((uintptr_t)p & 15u) == 0; };
  if (!aligned(src_lo) || !aligned(peer_dst_lo) || !aligned(src_hi) || !aligned(peer_dst_hi))
    return fail("fd_halo_push: planes must be 16-byte aligned");
  if ((src_lo == nullptr) != (peer_dst_lo == nullptr) || (src_hi == nullptr) != (peer_dst_hi == nullptr))
    return fail("fd_halo_push: a side needs both its source and its destination");
  cudaStream_t st = (cudaStream_t)stream;
  const long long n16 = nbytes / 16;
  const bool copy = n16 > 0 && (src_lo || src_hi), signal = peer_flag_lo || peer_flag_hi;
  if (!copy && !signal) return 0;
  if (copy) {
    long long blocks = (n16 + 255) / 256;
    if (blocks > 148 * 4) blocks = 148 * 4;
    halo_push_kernel<<<dim3((unsigned)blocks, 2), 256, 0, st>>>((const uint4*)src_lo, (uint4*)peer_dst_lo, (const uint4*)src_hi,
                                                                 (uint4*)peer_dst_hi, n16);
    if (signal) g_launches.fetch_add(1, std::memory_order_relaxed);
  }
  if (signal) halo_signal_kernel<<<1, 1, 0, st>>>((long long*)peer_flag_lo, (long long*)peer_flag_hi, value);
  return check_launch("fd_halo_push");
}

int fd_halo_wait(const void* flag_a, const void* flag_b, long long value, long long timeout_ns, void* flags, void* stream) {
  using namespace fdb;
  if (!flag_a && !flag_b) return 0;
  halo_wait_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((const long long*)flag_a, (const long long*)flag_b, value, timeout_ns,
                                                       (long long*)flags);
  return check_launch("fd_halo_wait");
}

int fd_halo_exchange(const fd_halo_link_t* links, int nlinks, void* counter, long long timeout_ns, void* flags, void* stream) {
  return fdb::launch_halo_exchange(links, nlinks, counter, timeout_ns, flags, (cudaStream_t)stream);
}

int fd_seq_create(void** seq_out) {
  using namespace fdb;
  if (!seq_out) return fail("fd_seq_create: NULL argument");
  Seq* q = new Seq();
  q->halo_recorded = false;
  q->counter = nullptr;
  cudaError_t e = cudaEventCreateWithFlags(&q->start, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&q->rims, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&q->halo, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&q->side_start, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&q->side_done, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc((void**)&q->counter, sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMemset(q->counter, 0, sizeof(unsigned));
  if (e != cudaSuccess) {
    cudaGetLastError();
    delete q;
    return fail("fd_seq_create: %s", cudaGetErrorString(e));
  }
  *seq_out = q;
  return 0;
}

int fd_seq_destroy(void* seq) {
  using namespace fdb;
  if (!seq) return 0;
  Seq* q = (Seq*)seq;
  cudaEventDestroy(q->start);
  cudaEventDestroy(q->rims);
  cudaEventDestroy(q->halo);
  cudaEventDestroy(q->side_start);
  cudaEventDestroy(q->side_done);
  cudaFree(q->counter);
  delete q;
  return 0;
}

int fd_seq_wait_halo(void* seq, void* stream) {
  using namespace fdb;
  if (!seq) return fail("fd_seq_wait_halo: NULL argument");
  Seq* q = (Seq*)seq;
  if (!q->halo_recorded) return 0;
  cudaError_t e = cudaStreamWaitEvent((cudaStream_t)stream, q->halo, 0);
  if (e != cudaSuccess) return fail("fd_seq_wait_halo: %s", cudaGetErrorString(e));
  return 0;
}

// one statement as several boxes (interior box + rim slabs of a fixed-boundary
// problem): the first box on `stream`, the others beside it on `side_stream`
static int step_boxes(bool f64, void* seq, const fd_fused_t* desc, int nboxes, const int32_t* boxes, void* stream, void* side_stream) {
  using namespace fdb;
  if (!seq || !desc || !boxes || nboxes < 1) return fail("fd_fused_step_boxes: bad argument");
  Seq* q = (Seq*)seq;
  cudaStream_t base = (cudaStream_t)stream, side = (cudaStream_t)side_stream;
  const bool beside = side_stream != nullptr && nboxes > 1;
  if (beside) {
    FD_CUDA(cudaEventRecord(q->side_start, base), "fd_fused_step_boxes");
    FD_CUDA(cudaStreamWaitEvent(side, q->side_start, 0), "fd_fused_step_boxes");
  }
  for (int i = 0; i < nboxes; ++i) {
    fd_fused_t d = *desc;
    for (int a = 0; a < 3; ++a) {
      d.box_lo[a] = boxes[6 * i + 2 * a];
      d.box_hi[a] = boxes[6 * i + 2 * a + 1];
    }
    void* where = (beside && i > 0) ? side_stream : stream;
    const int rc = f64 ? fd_fused_step_f64(&d, where) : fd_fused_step_f32(&d, where);
    if (rc) return rc;
  }
  if (beside) {
    FD_CUDA(cudaEventRecord(q->side_done, side), "fd_fused_step_boxes");
    FD_CUDA(cudaStreamWaitEvent(base, q->side_done, 0), "fd_fused_step_boxes");
  }
  return 0;
}

int fd_fused_step_boxes_f64(void* seq, const fd_fused_t* desc, int nboxes, const int32_t* boxes, void* stream, void* side_stream) {
  return step_boxes(true, seq, desc, nboxes, boxes, stream, side_stream);
}
int fd_fused_step_boxes_f32(void* seq, const fd_fused_t* desc, int nboxes, const int32_t* boxes, void* stream, void* side_stream) {
  return step_boxes(false, seq, desc, nboxes, boxes, stream, side_stream);
}

int fd_slab_step_f64(void* seq, const fd_slab_step_t* step) { return fdb::slab_step<double>(seq, step); }
int fd_slab_step_f32(void* seq, const fd_slab_step_t* step) { return fdb::slab_step<float>(seq, step); }

}  // extern "C"
