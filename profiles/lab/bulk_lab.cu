// Lab: 7-point fp64 periodic heat step with the planes staged in shared memory
// by cp.async.bulk (TMA bulk copies) + mbarrier pipeline, x marched in
// registers.  Compared bit-for-bit and in time against the register/L1 star
// kernel of the library on 512^3.  Not part of the library.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cstring>

#include "fd_star.cuh"

namespace fdb {
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
int fail(const char* fmt, ...) { fprintf(stderr, "fail: %s\n", fmt); return 1; }
}
using namespace fdb;

constexpr int TZ = 128, TY = 8, PZ = TZ + 4, ROWS = TY + 2;
constexpr int STAGE_DOUBLES = ROWS * PZ;
constexpr int STAGE_BYTES = STAGE_DOUBLES * 8;

struct BulkParams {
  const double* src;
  double* out;
  int n0, n1, n2, chunk;
  double wx[3], wy[3], wz[3];
  double mul, dt;
  long long* flags;
  long long* nan_count;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, int parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE;\nbra WAIT;\nDONE:\n}\n" ::"r"(
          smem_u32(b)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, int bytes, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(b))
               : "memory");
}

template <int NST>
__global__ void __launch_bounds__(256, 3) bulk_kernel(const __grid_constant__ BulkParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* stage = reinterpret_cast<double*>(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + NST * STAGE_BYTES);
  uint64_t* empty = full + NST;
  const int lane = threadIdx.x, row = threadIdx.y;
  const int tid = row * 32 + lane;
  const int z0 = blockIdx.x * TZ, y0 = blockIdx.y * TY;
  const int x_begin = blockIdx.z * P.chunk, x_end = min(x_begin + P.chunk, P.n0);
  const int nplanes = x_end - x_begin + 2;   // planes x_begin-1 .. x_end
  const int64_t plane = (int64_t)P.n1 * P.n2;

  if (tid == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // producer = warp 0: lane 0 arms the barrier, lanes 0..ROWS-1 copy one row each
  auto issue = [&](int q) {   // plane index q (0-based from x_begin-1) into stage q % NST
    int x = x_begin - 1 + q;
    x += (x < 0) ? P.n0 : 0;
    x -= (x >= P.n0) ? P.n0 : 0;
    const int s = q % NST;
    double* dst = stage + s * STAGE_DOUBLES;
    if (lane == 0) mbar_expect_tx(&full[s], STAGE_BYTES);
    __syncwarp();
    const double* base = P.src + x * plane;
    const bool left = z0 == 0, right = z0 + TZ == P.n2;
    if (lane < ROWS) {
      const int r = lane;
      int y = y0 - 1 + r;
      y += (y < 0) ? P.n1 : 0;
      y -= (y >= P.n1) ? P.n1 : 0;
      const double* rowp = base + (int64_t)y * P.n2;
      double* d = dst + r * PZ;
      if (!left && !right) {
        bulk_g2s(d, rowp + z0 - 2, PZ * 8, &full[s]);
      } else {
        // halo columns wrap around the periodic axis
        bulk_g2s(d, left ? rowp + P.n2 - 2 : rowp + z0 - 2, 16, &full[s]);
        bulk_g2s(d + 2, rowp + z0, TZ * 8, &full[s]);
        bulk_g2s(d + 2 + TZ, right ? rowp : rowp + z0 + TZ, 16, &full[s]);
      }
    }
  };

  if (row == 0)
    for (int q = 0; q < NST && q < nplanes; ++q) issue(q);

  unsigned nan_cnt = 0;
  bool saw_inf = false;
  // per-thread smem offsets (doubles): row (row+1), two vectors
  const int o0 = (row + 1) * PZ + 2 + lane * 2;
  const int o1 = o0 + 64;
  double2 wm[2], wc[2];   // x window: planes x-1, x
  {
    mbar_wait(&full[0], 0);
    const double* s0 = stage;
    wm[0] = *reinterpret_cast<const double2*>(s0 + o0);
    wm[1] = *reinterpret_cast<const double2*>(s0 + o1);
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[0]);
    mbar_wait(&full[1 % NST], 0);
    const double* s1 = stage + (1 % NST) * STAGE_DOUBLES;
    wc[0] = *reinterpret_cast<const double2*>(s1 + o0);
    wc[1] = *reinterpret_cast<const double2*>(s1 + o1);
  }
  double* outp = P.out + (int64_t)x_begin * plane + (int64_t)(y0 + row) * P.n2 + z0 + lane * 2;

  for (int i = 0; i < x_end - x_begin; ++i) {
    const int qc = i + 1, qn = i + 2;
    // producer: the stage of plane q=i is free once all warps released it
    if (row == 0 && i >= 1 && i - 1 + NST < nplanes) {
      if (lane == 0) mbar_wait(&empty[(i - 1) % NST], ((i - 1) / NST) & 1);
      __syncwarp();
      issue(i - 1 + NST);
    }
    mbar_wait(&full[qn % NST], (qn / NST) & 1);
    const double* sc = stage + (qc % NST) * STAGE_DOUBLES;
    const double* sn = stage + (qn % NST) * STAGE_DOUBLES;
    double2 wp[2];
    wp[0] = *reinterpret_cast<const double2*>(sn + o0);
    wp[1] = *reinterpret_cast<const double2*>(sn + o1);
    double2 res[2];
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      const int o = v ? o1 : o0;
      const double2 ym = *reinterpret_cast<const double2*>(sc + o - PZ);
      const double2 yp = *reinterpret_cast<const double2*>(sc + o + PZ);
      const double zl = sc[o - 1], zr = sc[o + 2];
      const double c[2] = {wc[v].x, wc[v].y};
      const double zw[4] = {zl, c[0], c[1], zr};
      const double ymv[2] = {ym.x, ym.y}, ypv[2] = {yp.x, yp.y};
      const double xm[2] = {wm[v].x, wm[v].y}, xp[2] = {wp[v].x, wp[v].y};
      double r[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        double dz = add_rn(add_rn(add_rn(0.0, mul_rn(P.wz[0], zw[e])), mul_rn(P.wz[1], zw[e + 1])), mul_rn(P.wz[2], zw[e + 2]));
        double dy = add_rn(add_rn(add_rn(0.0, mul_rn(P.wy[0], ymv[e])), mul_rn(P.wy[1], c[e])), mul_rn(P.wy[2], ypv[e]));
        double dx = add_rn(add_rn(add_rn(0.0, mul_rn(P.wx[0], xm[e])), mul_rn(P.wx[1], c[e])), mul_rn(P.wx[2], xp[e]));
        saw_inf |= is_inf(dz) | is_inf(dy) | is_inf(dx);
        double tot = mul_rn(add_rn(add_rn(dx, dy), dz), P.mul);
        double nv = add_rn(c[e], mul_rn(tot, P.dt));
        nan_cnt += is_nan(nv) ? 1u : 0u;
        r[e] = nv;
      }
      res[v] = make_double2(r[0], r[1]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[qc % NST]);   // plane qc no longer needed by this warp
    *reinterpret_cast<double2*>(outp) = res[0];
    *reinterpret_cast<double2*>(outp + 64) = res[1];
    outp += plane;
    wm[0] = wc[0]; wm[1] = wc[1];
    wc[0] = wp[0]; wc[1] = wp[1];
  }
  if (saw_inf) raise_flag(P.flags, FD_FLAG_INF);
  warp_add_count(P.nan_count, nan_cnt);
}


// Persistent variant: a fixed grid of blocks walks over the work items
// (z tile, y tile, x chunk) and streams ALL their planes through one smem ring,
// so the planes of the next item are already in flight while the current one is
// being finished (no per-item prologue bubble, no tail of partial waves).
template <int NST>
__global__ void __launch_bounds__(288, 3) bulk_persist_kernel(const __grid_constant__ BulkParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* stage = reinterpret_cast<double*>(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + NST * STAGE_BYTES);
  uint64_t* empty = full + NST;
  const int lane = threadIdx.x, row = threadIdx.y;
  const int tid = row * 32 + lane;
  const int64_t plane = (int64_t)P.n1 * P.n2;
  const int tiles_z = P.n2 / TZ, tiles_y = P.n1 / TY, chunks = P.n0 / P.chunk;
  const int nitems = tiles_z * tiles_y * chunks;
  const int PL = P.chunk + 2;                                   // planes streamed per item
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int total = my_items * PL;                              // planes this block streams

  if (tid == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto item_coords = [&](int k, int& z0, int& y0, int& x0) {
    const int item = blockIdx.x + k * gridDim.x;
    z0 = (item % tiles_z) * TZ;
    y0 = ((item / tiles_z) % tiles_y) * TY;
    x0 = (item / (tiles_z * tiles_y)) * P.chunk;
  };
  auto issue = [&](int g) {   // warp 0: plane g of this block's stream into stage g % NST
    int z0, y0, x0;
    item_coords(g / PL, z0, y0, x0);
    int x = x0 - 1 + g % PL;
    x += (x < 0) ? P.n0 : 0;
    x -= (x >= P.n0) ? P.n0 : 0;
    const int s = g % NST;
    double* dst = stage + s * STAGE_DOUBLES;
    if (lane == 0) mbar_expect_tx(&full[s], STAGE_BYTES);
    __syncwarp();
    const double* base = P.src + x * plane;
    const bool left = z0 == 0, right = z0 + TZ == P.n2;
    if (lane < ROWS) {
      int y = y0 - 1 + lane;
      y += (y < 0) ? P.n1 : 0;
      y -= (y >= P.n1) ? P.n1 : 0;
      const double* rowp = base + (int64_t)y * P.n2;
      double* d = dst + lane * PZ;
      if (!left && !right) {
        bulk_g2s(d, rowp + z0 - 2, PZ * 8, &full[s]);
      } else {
        bulk_g2s(d, left ? rowp + P.n2 - 2 : rowp + z0 - 2, 16, &full[s]);
        bulk_g2s(d + 2, rowp + z0, TZ * 8, &full[s]);
        bulk_g2s(d + 2 + TZ, right ? rowp : rowp + z0 + TZ, 16, &full[s]);
      }
    }
  };

  if (row == 8) {
    // dedicated producer warp: never delays the consumers' arithmetic
    for (int g = 0; g < total; ++g) {
      if (g >= NST) {
        if (lane == 0) mbar_wait(&empty[g % NST], ((g - NST) / NST) & 1);
        __syncwarp();
      }
      issue(g);
    }
    return;
  }

  unsigned nan_cnt = 0;
  bool saw_inf = false;
  const int o0 = (row + 1) * PZ + 2 + lane * 2;
  const int o1 = o0 + 64;
  double2 wm[2], wc[2];
  double* outp = nullptr;

  for (int g = 0; g < total; ++g) {
    const int q = g % PL;
    mbar_wait(&full[g % NST], (g / NST) & 1);
    const double* sn = stage + (g % NST) * STAGE_DOUBLES;
    double2 wp[2];
    wp[0] = *reinterpret_cast<const double2*>(sn + o0);
    wp[1] = *reinterpret_cast<const double2*>(sn + o1);
    if (q == 0) {
      int z0, y0, x0;
      item_coords(g / PL, z0, y0, x0);
      outp = P.out + (int64_t)x0 * plane + (int64_t)(y0 + row) * P.n2 + z0 + lane * 2;
      wc[0] = wp[0]; wc[1] = wp[1];
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[g % NST]);
      continue;
    }
    if (q == 1) {
      wm[0] = wc[0]; wm[1] = wc[1];
      wc[0] = wp[0]; wc[1] = wp[1];
      continue;
    }
    const double* sc = stage + ((g - 1) % NST) * STAGE_DOUBLES;
    double2 res[2];
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      const int o = v ? o1 : o0;
      const double2 ym = *reinterpret_cast<const double2*>(sc + o - PZ);
      const double2 yp = *reinterpret_cast<const double2*>(sc + o + PZ);
      const double zl = sc[o - 1], zr = sc[o + 2];
      const double c[2] = {wc[v].x, wc[v].y};
      const double zw[4] = {zl, c[0], c[1], zr};
      const double ymv[2] = {ym.x, ym.y}, ypv[2] = {yp.x, yp.y};
      const double xm[2] = {wm[v].x, wm[v].y}, xp[2] = {wp[v].x, wp[v].y};
      double r[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        double dz = add_rn(add_rn(add_rn(0.0, mul_rn(P.wz[0], zw[e])), mul_rn(P.wz[1], zw[e + 1])), mul_rn(P.wz[2], zw[e + 2]));
        double dy = add_rn(add_rn(add_rn(0.0, mul_rn(P.wy[0], ymv[e])), mul_rn(P.wy[1], c[e])), mul_rn(P.wy[2], ypv[e]));
        double dx = add_rn(add_rn(add_rn(0.0, mul_rn(P.wx[0], xm[e])), mul_rn(P.wx[1], c[e])), mul_rn(P.wx[2], xp[e]));
        saw_inf |= is_inf(dz) | is_inf(dy) | is_inf(dx);
        double tot = mul_rn(add_rn(add_rn(dx, dy), dz), P.mul);
        double nv = add_rn(c[e], mul_rn(tot, P.dt));
        nan_cnt += is_nan(nv) ? 1u : 0u;
        r[e] = nv;
      }
      res[v] = make_double2(r[0], r[1]);
    }
    __syncwarp();
    if (lane == 0) {
      mbar_arrive(&empty[(g - 1) % NST]);            // the centre plane is done
      if (q == PL - 1) mbar_arrive(&empty[g % NST]);  // the last plane of an item is never a centre
    }
    *reinterpret_cast<double2*>(outp) = res[0];
    *reinterpret_cast<double2*>(outp + 64) = res[1];
    outp += plane;
    wm[0] = wc[0]; wm[1] = wc[1];
    wc[0] = wp[0]; wc[1] = wp[1];
  }
  if (saw_inf) raise_flag(P.flags, FD_FLAG_INF);
  warp_add_count(P.nan_count, nan_cnt);
}

int main() {
  const int n = 512;
  const size_t N = (size_t)n * n * n;
  double *a, *b, *c;
  long long* flags;
  cudaMalloc(&a, N * 8); cudaMalloc(&b, N * 8); cudaMalloc(&c, N * 8);
  cudaMalloc(&flags, 64); cudaMemset(flags, 0, 64);
  std::vector<double> h(N);
  for (size_t i = 0; i < N; ++i) h[i] = (double)rand() / RAND_MAX;
  cudaMemcpy(a, h.data(), N * 8, cudaMemcpyHostToDevice);

  // reference result with the library's star kernel
  fd_fused_t d = {};
  d.n[0] = d.n[1] = d.n[2] = n;
  d.nterms = 3;
  for (int k = 0; k < 3; ++k) {
    fd_term_t& t = d.term[k];
    t.src = a; t.axis = k; t.periodic = 1; t.lo = 0; t.hi = n; t.n_src = n;
    t.st.npts = 3;
    for (int p = 0; p < 3; ++p) t.st.shifts[p] = p - 1;
    t.st.weights[0] = 1; t.st.weights[1] = -2; t.st.weights[2] = 1;
    t.st.scale = 1.0; t.st.exact_div = 1;
  }
  d.npost = 1; d.post[0].op = FD_POST_MUL; d.post[0].s = 0.1;
  d.depth = 1; d.dt = 1e-3;
  d.prev[0] = a; d.out[0] = b; d.flags = flags;
  for (int k = 0; k < 3; ++k) { d.rim_lo[k] = 0; d.rim_hi[k] = n - 1; }
  DevFused<double> F;
  if (make_dev_fused<double>(&d, &F)) return 1;
  StarParams<double> SP;
  int R, axes, mode;
  if (!star_applicable<double>(F, &SP, &R, &axes, &mode)) { printf("not applicable\n"); return 1; }
  SP.x_chunk = 8;
  {
    dim3 grid(n / 2 / 64, n / 8, n / 8), block(32, 8, 1);
    star_kernel<double, 1, 3, 0, 2, 3><<<grid, block>>>(SP);
  }
  cudaDeviceSynchronize();

  BulkParams P = {};
  P.src = a; P.out = c; P.n0 = P.n1 = P.n2 = n;
  for (int p = 0; p < 3; ++p) P.wx[p] = P.wy[p] = P.wz[p] = (p == 1 ? -2.0 : 1.0);
  P.mul = 0.1; P.dt = 1e-3; P.flags = flags; P.nan_count = flags + 2;
  auto run = [&](auto kern, int nst, int chunk, int iters) {
    const int smem = nst * STAGE_BYTES + 2 * nst * 8;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    P.chunk = chunk;
    dim3 grid(n / TZ, n / TY, (n + chunk - 1) / chunk), block(32, 8, 1);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int it = 0; it < 3 + iters; ++it) {
      if (it == 3) cudaEventRecord(e0);
      kern<<<grid, block, smem>>>(P);
    }
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(err));
    return ms / iters;
  };
  float t = run(bulk_kernel<4>, 4, 8, 1);
  std::vector<double> hb(N), hc(N);
  cudaMemcpy(hb.data(), b, N * 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(hc.data(), c, N * 8, cudaMemcpyDeviceToHost);
  size_t bad = 0;
  for (size_t i = 0; i < N; ++i) bad += memcmp(&hb[i], &hc[i], 8) != 0;
  printf("mismatches vs star kernel: %zu of %zu\n", bad, N);
  auto runp = [&](auto kern, int nst, int chunk, int blocks_per_sm, int iters) {
    const int smem = nst * STAGE_BYTES + 2 * nst * 8;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    P.chunk = chunk;
    dim3 grid(148 * blocks_per_sm), block(32, 9, 1);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int it = 0; it < 3 + iters; ++it) {
      if (it == 3) cudaEventRecord(e0);
      kern<<<grid, block, smem>>>(P);
    }
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(err));
    return ms / iters;
  };
  cudaMemset(c, 0, N * 8);
  runp(bulk_persist_kernel<6>, 6, 16, 3, 1);
  cudaMemcpy(hc.data(), c, N * 8, cudaMemcpyDeviceToHost);
  bad = 0;
  for (size_t i = 0; i < N; ++i) bad += memcmp(&hb[i], &hc[i], 8) != 0;
  printf("persistent: mismatches vs star kernel: %zu of %zu\n", bad, N);
  for (int chunk : {8, 16, 32, 64})
    printf("persistent chunk %2d: nst4x3 %.4f  nst6x3 %.4f  nst6x2 %.4f  nst8x2 %.4f ms\n", chunk,
           runp(bulk_persist_kernel<4>, 4, chunk, 3, 20), runp(bulk_persist_kernel<6>, 6, chunk, 3, 20),
           runp(bulk_persist_kernel<6>, 6, chunk, 2, 20), runp(bulk_persist_kernel<8>, 8, chunk, 2, 20));
  for (int chunk : {16}) {
    printf("chunk %2d: nst3 %.4f  nst4 %.4f  nst6 %.4f ms\n", chunk, run(bulk_kernel<3>, 3, chunk, 20),
           run(bulk_kernel<4>, 4, chunk, 20), run(bulk_kernel<6>, 6, chunk, 20));
  }
  return 0;
}
