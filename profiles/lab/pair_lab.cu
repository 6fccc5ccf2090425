// Tuning / parity lab for the staggered pair kernels, through the C ABI of the
// built library:  pair_lab nx ny nz [f32|f64] [spacing]
// Runs fd_fused_pair with FD_B200_NO_MARCH=1 (pair_kernel) and without (marching
// kernel) on the same random state, compares the four outputs and the NaN
// counters bit for bit (also with NaN / inf planted in the inputs), and times a
// few tuning variants.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "fd_b200.h"

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e_ = (x);                                                             \
    if (e_ != cudaSuccess) {                                                          \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                        \
    }                                                                                 \
  } while (0)

template <typename T>
__global__ void fill_random(T* a, size_t n, unsigned seed) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    unsigned h = (unsigned)(i * 2654435761u) ^ seed;
    h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
    a[i] = (T)((h & 0xffffff) / (double)0x1000000 - 0.5);
  }
}

template <typename T>
__global__ void plant(T* a, size_t n, size_t stride, T value) {
  for (size_t i = stride / 3 + (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * stride; i < n; i += (size_t)gridDim.x * blockDim.x * stride)
    a[i] = value;
}

__global__ void count_diff(const unsigned* a, const unsigned* b, size_t nwords, unsigned long long* out) {
  unsigned long long c = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x)
    c += a[i] != b[i];
  if (c) atomicAdd(out, c);
}

template <typename T>
int run(int nx, int ny, int nz, double spacing, bool slab) {
  const size_t N = (size_t)nx * ny * nz;
  T *in[4], *out_a[4], *out_b[4];
  for (int k = 0; k < 4; ++k) {
    CK(cudaMalloc(&in[k], N * sizeof(T)));
    CK(cudaMalloc(&out_a[k], N * sizeof(T)));
    CK(cudaMalloc(&out_b[k], N * sizeof(T)));
  }
  long long* counters;   // [0..3] flags run a, [4..7] flags run b, [8..11] NaN counters run a, [16..19] run b
  CK(cudaMalloc(&counters, 32 * sizeof(long long)));
  unsigned long long* diff;
  CK(cudaMalloc(&diff, 8));

  auto describe = [&](T** out, long long* nan, long long* flags) {
    fd_pair_t d;
    memset(&d, 0, sizeof d);
    d.n[0] = nx; d.n[1] = ny; d.n[2] = nz;
    d.naxes = 3;
    for (int a = 0; a < 3; ++a) {
      fd_pair_axis_t& A = d.ax[a];
      A.v_prev = in[1 + a]; A.v_out = out[1 + a]; A.v_nan_count = nan + 1 + a;
      A.c_up = -1.0; A.axis = a; A.periodic = 1; A.subtract = a == 2;
      A.up.npts = 2; A.up.shifts[0] = -1; A.up.shifts[1] = 0; A.up.weights[0] = -1; A.up.weights[1] = 1;
      A.up.scale = spacing; A.up.exact_div = 1;
      A.down.npts = 2; A.down.shifts[0] = 0; A.down.shifts[1] = 1; A.down.weights[0] = -1; A.down.weights[1] = 1;
      A.down.scale = spacing; A.down.exact_div = 1;
    }
    d.p_prev = in[0]; d.p_out = out[0]; d.p_nan_count = nan;
    d.flags = flags;
    d.c_down = -0.25; d.dt = 1e-3;
    if (slab) {   // x is the decomposed axis of a slab: one ghost plane on each side, owned planes 1 .. nx-2
      d.ax[0].periodic = 0;
      d.box_lo[0] = 1; d.box_hi[0] = nx - 1;
      d.box_hi[1] = ny; d.box_hi[2] = nz;
    }
    return d;
  };
  auto call = [&](const fd_pair_t& d) {
    const int rc = sizeof(T) == 4 ? fd_fused_pair_f32(&d, nullptr) : fd_fused_pair_f64(&d, nullptr);
    if (rc) { printf("fd_fused_pair failed: %s\n", fd_last_error()); exit(1); }
  };
  auto timed = [&](const fd_pair_t& d, int reps) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    call(d);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) call(d);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / reps;
  };

  fd_pair_t da = describe(out_a, counters + 8, counters), db = describe(out_b, counters + 16, counters + 4);
  int bad = 0;
  setenv("FD_B200_PAIR_CHUNK", "8", 1);            // a forced chunk also forces the marching kernel on small grids
  for (int round = 0; round < 3; ++round) {        // 0: finite data, 1: NaN planted, 2: inf planted
    for (int k = 0; k < 4; ++k) fill_random<T><<<1184, 256>>>(in[k], N, 77u + k + 10 * round);
    if (round == 1) { plant<T><<<64, 64>>>(in[0], N, 99991, (T)NAN); plant<T><<<64, 64>>>(in[2], N, 70001, (T)NAN); }
    if (round == 2) { plant<T><<<64, 64>>>(in[0], N, 99991, (T)INFINITY); plant<T><<<64, 64>>>(in[3], N, 50021, (T)-INFINITY); }
    CK(cudaMemset(counters, 0, 32 * sizeof(long long)));
    for (int vy = 2; vy >= 1; --vy) {
      char buf[8];
      snprintf(buf, sizeof buf, "%d", vy);
      setenv("FD_B200_PAIR_VY", buf, 1);
      for (int k = 0; k < 4; ++k) { CK(cudaMemset(out_b[k], 0xff, N * sizeof(T))); if (vy == 2) CK(cudaMemset(out_a[k], 0xff, N * sizeof(T))); }
      if (vy == 2) { setenv("FD_B200_NO_MARCH", "1", 1); CK(cudaMemset(counters, 0, 4 * sizeof(long long))); CK(cudaMemset(counters + 8, 0, 8 * sizeof(long long))); call(da); setenv("FD_B200_NO_MARCH", "0", 1); }
      CK(cudaMemset(counters + 4, 0, 4 * sizeof(long long)));
      CK(cudaMemset(counters + 16, 0, 8 * sizeof(long long)));
      call(db);
      CK(cudaDeviceSynchronize());
      for (int k = 0; k < 4; ++k) {
        CK(cudaMemset(diff, 0, 8));
        count_diff<<<1184, 256>>>((const unsigned*)out_a[k], (const unsigned*)out_b[k], N * sizeof(T) / 4, diff);
        unsigned long long h;
        CK(cudaMemcpy(&h, diff, 8, cudaMemcpyDeviceToHost));
        if (h) { printf("round %d VY=%d output %d: %llu words differ\n", round, vy, k, h); bad = 1; }
      }
      long long hc[32];
      CK(cudaMemcpy(hc, counters, sizeof hc, cudaMemcpyDeviceToHost));
      for (int k = 0; k < 4; ++k)
        if (hc[8 + k] != hc[16 + k]) { printf("round %d VY=%d NaN counter %d: %lld vs %lld\n", round, vy, k, hc[8 + k], hc[16 + k]); bad = 1; }
      for (int k = 0; k < 4; ++k)
        if ((hc[k] != 0) != (hc[4 + k] != 0)) { printf("round %d VY=%d flag %d: %lld vs %lld\n", round, vy, k, hc[k], hc[4 + k]); bad = 1; }
      if (vy == 1) printf("  round %d: NaN counts %lld %lld %lld %lld, inf flag %lld\n", round, hc[16], hc[17], hc[18], hc[19], hc[4]);
    }
  }
  printf("%s %dx%dx%d spacing %g%s: marching kernel %s pair_kernel\n", sizeof(T) == 4 ? "f32" : "f64", nx, ny, nz, spacing, slab ? " (slab box)" : "",
         bad ? "DIFFERS FROM" : "bit-identical to");

  for (int k = 0; k < 4; ++k) fill_random<T><<<1184, 256>>>(in[k], N, 77u + k);
  const double gb = 8.0 * N * sizeof(T) / 1e9;
  setenv("FD_B200_NO_MARCH", "1", 1);
  float t = timed(da, 5);
  printf("  pair_kernel                 %8.3f ms  %7.1f GB/s\n", t, gb / t * 1e3);
  setenv("FD_B200_NO_MARCH", "0", 1);
  for (int vy = 1; vy <= 2; ++vy)
    for (int chunk : {8, 16, 32, 64}) {
      char b1[8], b2[8];
      snprintf(b1, sizeof b1, "%d", vy);
      snprintf(b2, sizeof b2, "%d", chunk);
      setenv("FD_B200_PAIR_VY", b1, 1);
      setenv("FD_B200_PAIR_CHUNK", b2, 1);
      t = timed(db, 5);
      printf("  march VY=%d chunk=%-3d        %8.3f ms  %7.1f GB/s\n", vy, chunk, t, gb / t * 1e3);
    }
  unsetenv("FD_B200_PAIR_VY");
  unsetenv("FD_B200_PAIR_CHUNK");
  t = timed(db, 5);
  printf("  march (library defaults)    %8.3f ms  %7.1f GB/s\n", t, gb / t * 1e3);
  for (int k = 0; k < 4; ++k) { cudaFree(in[k]); cudaFree(out_a[k]); cudaFree(out_b[k]); }
  cudaFree(counters); cudaFree(diff);
  return bad;
}

int main(int argc, char** argv) {
  int nx = 256, ny = 2048, nz = 2048;
  if (argc > 3) { nx = atoi(argv[1]); ny = atoi(argv[2]); nz = atoi(argv[3]); }
  const bool f64 = argc > 4 && !strcmp(argv[4], "f64");
  const double spacing = argc > 5 ? atof(argv[5]) : 1.0;
  const bool slab = argc > 6 && !strcmp(argv[6], "slab");
  return f64 ? run<double>(nx, ny, nz, spacing, slab) : run<float>(nx, ny, nz, spacing, slab);
}
