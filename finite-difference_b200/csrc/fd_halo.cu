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

}  // extern "C"
