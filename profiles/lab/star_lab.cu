// Tuning lab for the star kernel on the 512^3 fp64 7-point periodic case
// (BASELINE.json configs[2]).  Not part of the library: built and run by
// profiles/lab/run_lab.sh on the GPU box; prints ms per launch for each
// launch configuration.  Input: random doubles; buffers ping-pong.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fd_star.cuh"

namespace fdb {
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
int fail(const char* fmt, ...) { fprintf(stderr, "fail: %s\n", fmt); return 1; }
}
using namespace fdb;

static int g_carveout = -1;

template <int OCC, int VPT>
static float run(StarParams<double> P, double* a, double* b, int bx, int by, int chunk, int iters) {
  const int n0 = P.F.n[0], n1 = P.F.n[1], nvec = P.F.n[2] / 2 / VPT;
  dim3 grid((P.F.n[2] / 2 + bx * VPT - 1) / (bx * VPT), (n1 + by - 1) / by, (n0 + chunk - 1) / chunk), block(bx, by, 1);
  P.x_chunk = chunk;
  cudaFuncSetAttribute(star_kernel<double, 1, 3, 0, VPT, OCC>, cudaFuncAttributePreferredSharedMemoryCarveout, g_carveout);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int it = 0; it < 3 + iters; ++it) {
    if (it == 3) cudaEventRecord(e0);
    P.src = a; P.F.prev[0] = a; P.F.out[0] = b;
    star_kernel<double, 1, 3, 0, VPT, OCC><<<grid, block>>>(P);
    std::swap(a, b);
  }
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(err));
  return ms / iters;
}

int main() {
  const int n = 512;
  const size_t N = (size_t)n * n * n;
  double *a, *b;
  long long* flags;
  cudaMalloc(&a, N * 8);
  cudaMalloc(&b, N * 8);
  cudaMalloc(&flags, 64);
  cudaMemset(flags, 0, 64);
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
  DevFused<double> F;
  if (make_dev_fused<double>(&d, &F)) return 1;
  StarParams<double> P;
  int R, axes, mode;
  if (!star_applicable<double>(F, &P, &R, &axes, &mode)) { printf("not applicable\n"); return 1; }
  printf("R=%d axes=%d mode=%d\n", R, axes, mode);
  const int iters = 30;
  for (int cv : {-1, 0, 25, 50, 100}) {
    g_carveout = cv;
    printf("carveout %3d: vpt2/occ3 32x8 chunk8 %.4f ms   chunk16 %.4f ms\n", cv, run<3, 2>(P, a, b, 32, 8, 8, 30),
           run<3, 2>(P, a, b, 32, 8, 16, 30));
  }
  return 0;
}
