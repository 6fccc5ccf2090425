// Lab: thread-to-point mappings for the marched 7-point fp64 periodic step
// (R = 1): VY rows x VZ interleaved 128-bit vectors per thread.  Same
// arithmetic as the library's star kernel (checked bit-for-bit against it).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "fd_star.cuh"

namespace fdb {
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
int fail(const char* fmt, ...) { fprintf(stderr, "fail: %s\n", fmt); return 1; }
}
using namespace fdb;

struct MapParams {
  const double* src;
  double* out;
  int n0, n1, n2, chunk;
  double wx[3], wy[3], wz[3];
  double mul, dt;
  long long* flags;
  long long* nan_count;
};

__device__ __forceinline__ double2 ldg2(const double* p) { return __ldg(reinterpret_cast<const double2*>(p)); }

template <int VY, int VZ, int OCC>
__global__ void __launch_bounds__(256, OCC) map_kernel(const __grid_constant__ MapParams P) {
  const int n0 = P.n0, n1 = P.n1, n2 = P.n2;
  const int nvec = n2 / 2;
  const int vz0 = blockIdx.x * (32 * VZ) + threadIdx.x;            // first vector; others +32
  const int y0 = (blockIdx.y * blockDim.y + threadIdx.y) * VY;     // first row; others +1
  const int x_begin = blockIdx.z * P.chunk, x_end = min(x_begin + P.chunk, n0);
  const int plane = n1 * n2;
  unsigned nan_cnt = 0;
  bool saw_inf = false;
  // offsets (elements) relative to (x, y0, vz0*2)
  int zl[VZ], zr[VZ];
#pragma unroll
  for (int v = 0; v < VZ; ++v) {
    int a = vz0 + 32 * v - 1, b = vz0 + 32 * v + 1;
    a += a < 0 ? nvec : 0;
    b -= b >= nvec ? nvec : 0;
    zl[v] = (a - vz0) * 2 + 1;        // the element just left of the vector
    zr[v] = (b - vz0) * 2;            // the element just right of it
  }
  int ym = y0 - 1, yp = y0 + VY;
  ym += ym < 0 ? n1 : 0;
  yp -= yp >= n1 ? n1 : 0;
  const int oym = (ym - y0) * n2, oyp = (yp - y0) * n2;
  const double* col = P.src + (int64_t)y0 * n2 + vz0 * 2;
  auto wrapx = [&](int x) { x += x < 0 ? n0 : 0; x -= x >= n0 ? n0 : 0; return x; };
  double2 wm[VY][VZ], wc[VY][VZ], wp[VY][VZ], wq[VY][VZ];   // planes x-1, x, x+1 and the prefetched x+2
  auto load_plane = [&](double2 (&w)[VY][VZ], int x) {
    const double* p = col + (int64_t)wrapx(x) * plane;
#pragma unroll
    for (int r = 0; r < VY; ++r)
#pragma unroll
      for (int v = 0; v < VZ; ++v) w[r][v] = ldg2(p + r * n2 + 64 * v);
  };
  load_plane(wm, x_begin - 1);
  load_plane(wc, x_begin);
  load_plane(wp, x_begin + 1);
  const double* here = col + (int64_t)x_begin * plane;
  double* outp = P.out + (int64_t)x_begin * plane + (int64_t)y0 * n2 + vz0 * 2;
  for (int x = x_begin; x < x_end; ++x) {
    load_plane(wq, x + 2);
    double2 yrow_m[VZ], yrow_p[VZ];
#pragma unroll
    for (int v = 0; v < VZ; ++v) {
      yrow_m[v] = ldg2(here + oym + 64 * v);
      yrow_p[v] = ldg2(here + oyp + 64 * v);
    }
#pragma unroll
    for (int r = 0; r < VY; ++r) {
#pragma unroll
      for (int v = 0; v < VZ; ++v) {
        const double zlv = __ldg(here + r * n2 + zl[v]), zrv = __ldg(here + r * n2 + zr[v]);
        const double2 ym2 = r == 0 ? yrow_m[v] : wc[r - 1 < 0 ? 0 : r - 1][v];
        const double2 yp2 = r == VY - 1 ? yrow_p[v] : wc[r + 1 >= VY ? VY - 1 : r + 1][v];
        const double c[2] = {wc[r][v].x, wc[r][v].y};
        const double zw[4] = {zlv, c[0], c[1], zrv};
        const double ymv[2] = {ym2.x, ym2.y}, ypv[2] = {yp2.x, yp2.y};
        const double xm[2] = {wm[r][v].x, wm[r][v].y}, xp[2] = {wp[r][v].x, wp[r][v].y};
        double res[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          double dz = add_rn(add_rn(add_rn(0.0, mul_rn(P.wz[0], zw[e])), mul_rn(P.wz[1], zw[e + 1])), mul_rn(P.wz[2], zw[e + 2]));
          double dy = add_rn(add_rn(add_rn(0.0, mul_rn(P.wy[0], ymv[e])), mul_rn(P.wy[1], c[e])), mul_rn(P.wy[2], ypv[e]));
          double dx = add_rn(add_rn(add_rn(0.0, mul_rn(P.wx[0], xm[e])), mul_rn(P.wx[1], c[e])), mul_rn(P.wx[2], xp[e]));
          saw_inf |= is_inf(dz) | is_inf(dy) | is_inf(dx);
          double tot = mul_rn(add_rn(add_rn(dx, dy), dz), P.mul);
          double nv = add_rn(c[e], mul_rn(tot, P.dt));
          nan_cnt += is_nan(nv) ? 1u : 0u;
          res[e] = nv;
        }
        *reinterpret_cast<double2*>(outp + r * n2 + 64 * v) = make_double2(res[0], res[1]);
      }
    }
    here += plane;
    outp += plane;
#pragma unroll
    for (int r = 0; r < VY; ++r)
#pragma unroll
      for (int v = 0; v < VZ; ++v) {
        wm[r][v] = wc[r][v];
        wc[r][v] = wp[r][v];
        wp[r][v] = wq[r][v];
      }
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
  if (!star_applicable<double>(F, &SP, &R, &axes, &mode)) return 1;
  SP.x_chunk = 8;
  {
    dim3 grid(n / 2 / 64, n / 8, n / 8), block(32, 8, 1);
    star_kernel<double, 1, 3, 0, 2, 3><<<grid, block>>>(SP);
  }
  cudaDeviceSynchronize();
  std::vector<double> hb(N), hc(N);
  cudaMemcpy(hb.data(), b, N * 8, cudaMemcpyDeviceToHost);

  MapParams P = {};
  P.src = a; P.out = c; P.n0 = P.n1 = P.n2 = n;
  for (int p = 0; p < 3; ++p) P.wx[p] = P.wy[p] = P.wz[p] = (p == 1 ? -2.0 : 1.0);
  P.mul = 0.1; P.dt = 1e-3; P.flags = flags; P.nan_count = flags + 2;
  auto run = [&](auto kern, int vy, int vz, int by, int chunk, int iters, bool check) {
    P.chunk = chunk;
    dim3 grid(n / 2 / (32 * vz), n / (by * vy), (n + chunk - 1) / chunk), block(32, by, 1);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int it = 0; it < 3 + iters; ++it) {
      if (it == 3) cudaEventRecord(e0);
      kern<<<grid, block>>>(P);
    }
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(err));
    if (check) {
      cudaMemcpy(hc.data(), c, N * 8, cudaMemcpyDeviceToHost);
      size_t bad = 0;
      for (size_t i = 0; i < N; ++i) bad += memcmp(&hb[i], &hc[i], 8) != 0;
      printf("   mismatches: %zu\n", bad);
    }
    return ms / iters;
  };
  for (int chunk : {4, 6, 8, 12}) {
    printf("chunk %d\n", chunk);
    printf(" VY2 VZ1 occ4 by8 : %.4f ms\n", run(map_kernel<2, 1, 4>, 2, 1, 8, chunk, 20, chunk == 8));
    printf(" VY2 VZ1 occ4 by4 : %.4f ms\n", run(map_kernel<2, 1, 4>, 2, 1, 4, chunk, 20, false));
    printf(" VY2 VZ1 occ4 by2 : %.4f ms\n", run(map_kernel<2, 1, 4>, 2, 1, 2, chunk, 20, false));
    printf(" VY2 VZ1 occ5 by8 : %.4f ms\n", run(map_kernel<2, 1, 5>, 2, 1, 8, chunk, 20, false));
    printf(" VY1 VZ2 occ4 by8 : %.4f ms\n", run(map_kernel<1, 2, 4>, 1, 2, 8, chunk, 20, false));
    printf(" VY3 VZ1 occ3 by8 : %.4f ms\n", run(map_kernel<3, 1, 3>, 3, 1, 8, chunk, 20, false));
    printf(" VY2 VZ2 occ2 by4 : %.4f ms\n", run(map_kernel<2, 2, 2>, 2, 2, 4, chunk, 20, false));
  }
  return 0;
}
