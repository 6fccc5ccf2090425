// One configuration of the star kernel on 512^3 (for ncu): star_one <bx> <by> <chunk>
#define main lab_main
#include "star_lab.cu"
#undef main
int main(int argc, char** argv) {
  const int bx = argc > 1 ? atoi(argv[1]) : 32, by = argc > 2 ? atoi(argv[2]) : 8, chunk = argc > 3 ? atoi(argv[3]) : 16;
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
  make_dev_fused<double>(&d, &F);
  StarParams<double> P;
  int R, axes, mode;
  star_applicable<double>(F, &P, &R, &axes, &mode);
  printf("%.4f ms\n", run<3, 2>(P, a, b, bx, by, chunk, 10));
  return 0;
}
