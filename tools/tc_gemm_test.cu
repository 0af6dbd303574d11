// Standalone check of the tcgen05 GEMM engine (gpflow_b200/csrc/gpflow_tc_gemm.cuh) against an fp64 host reference.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o tools/tc_gemm_test tools/tc_gemm_test.cu
//   ./tools/tc_gemm_test            (prints max relative errors and a TF32x3 throughput figure)
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <type_traits>
#include <vector>

#define GPF_TC_PROF 1
#include "../gpflow_b200/csrc/gpflow_tc_gemm.cuh"

using namespace gpflow;
using namespace gpflow::tc;

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess) {                                                               \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);      \
      exit(2);                                                                             \
    }                                                                                      \
  } while (0)

// ---- op 1: D[M][N] = A[M][K] W[N][K]^T, A from global rows, W pre-split via TMA ----
template <int NT_, int ST_>
struct OpPlain {
  static constexpr int NT = NT_, STAGES = ST_;
  static constexpr bool B_TMA = true;
  static constexpr int A_MODE = MODE_ROWS, B_MODE = MODE_ROWS;
  struct Params {
    CUtensorMap map;   // over Wsplit [2 * npad][K]
    const float* A;
    float* D;
    long long* prof;
    int dbg;   // 1: skip A production, 2: skip TMA loads, 4: skip epilogue stores
    int M, N, K, npad, mtiles, ntiles_n;
  };
  struct Tile { int nk, n, m0, n0; };
  GPF_DEV static int num_tiles(const Params& p) { return p.mtiles * p.ntiles_n; }
  GPF_DEV static void decode(const Params& p, int t, Tile& tl) {
    tl.m0 = (t / p.ntiles_n) * MT;
    tl.n0 = (t % p.ntiles_n) * NT;
    tl.nk = (p.K + KC - 1) / KC;
    const int rem = p.N - tl.n0;
    tl.n = rem >= NT ? NT : (rem + 15) / 16 * 16;
  }
  GPF_DEV static const CUtensorMap* tmap(const Params& p, const Tile&) { return &p.map; }
  GPF_DEV static void tma_coords(const Params& p, const Tile& tl, int kc, int& c0, int& r_hi, int& r_lo) {
    c0 = kc * KC; r_hi = tl.n0; r_lo = p.npad + tl.n0;
  }
  GPF_DEV static void fetch_a(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    const int k0 = kc * KC;
    if (p.dbg & 1) return;
    fetch_rows<MT>(frag, ptid, p.A, p.K, tl.m0, p.M, k0, p.K - k0);
  }
  GPF_DEV static void epilogue(const Params& p, const Tile& tl, const EpiCtx& ctx, int row, int c0, int nc, const float (&v)[32]) {
    if (p.dbg & 4) return;
    if ((p.N & 3) == 0 && tl.n0 + c0 + nc <= p.N) {   // full chunk: coalesced row stores through shared memory
      ctx.store_rows(v, nc, c0, [&](int r) -> float* {
        return tl.m0 + r < p.M ? p.D + (size_t)(tl.m0 + r) * p.N + tl.n0 : nullptr; }, false);
      return;
    }
    const int m = tl.m0 + row;
    if (m >= p.M) return;
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < nc && tl.n0 + c0 + i < p.N) p.D[(size_t)m * p.N + tl.n0 + c0 + i] = v[i];
  }
};

// ---- op 2: D[M][N] = sum_k X[k][m] Y[k][n]  (both operands transposed by the producers, no TMA) ----
template <int NT_, int ST_>
struct OpTrans {
  static constexpr int NT = NT_, STAGES = ST_;
  static constexpr bool B_TMA = false;
  static constexpr int A_MODE = MODE_UNITS, B_MODE = MODE_UNITS;
  struct Params {
    const float* X;   // [K][M]
    const float* Y;   // [K][N]
    float* D;
    long long* prof;
    int dbg;
    int M, N, K, mtiles, ntiles_n;
  };
  struct Tile { int nk, n, m0, n0; };
  GPF_DEV static int num_tiles(const Params& p) { return p.mtiles * p.ntiles_n; }
  GPF_DEV static void decode(const Params& p, int t, Tile& tl) {
    tl.m0 = (t / p.ntiles_n) * MT;
    tl.n0 = (t % p.ntiles_n) * NT;
    tl.nk = (p.K + KC - 1) / KC;
    const int rem = p.N - tl.n0;
    tl.n = rem >= NT ? NT : (rem + 15) / 16 * 16;
  }
  GPF_DEV static void fetch_a(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    fetch_units<MT>(frag, ptid, [&](int m, int kq, float (&v)[8]) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = kc * KC + kq * 8 + i, mm = tl.m0 + m;
        v[i] = (k < p.K && mm < p.M) ? p.X[(size_t)k * p.M + mm] : 0.f;
      }
    });
  }
  GPF_DEV static void fetch_b(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    fetch_units<NT>(frag, ptid, [&](int n, int kq, float (&v)[8]) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = kc * KC + kq * 8 + i, nn = tl.n0 + n;
        v[i] = (k < p.K && nn < p.N) ? p.Y[(size_t)k * p.N + nn] : 0.f;
      }
    });
  }
  GPF_DEV static void epilogue(const Params& p, const Tile& tl, const EpiCtx& ctx, int row, int c0, int nc, const float (&v)[32]) {
    const int m = tl.m0 + row;
    if (m >= p.M) return;
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < nc && tl.n0 + c0 + i < p.N) p.D[(size_t)m * p.N + tl.n0 + c0 + i] = v[i];
  }
};

__global__ void split_kernel(const float* w, float* out, int n_rows, int K, int npad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows * K) return;
  float hi, lo;
  split_tf32(w[i], hi, lo);
  out[i] = hi;
  out[(size_t)npad * K + i] = lo;
}

static double frand() { return (double)rand() / RAND_MAX * 2.0 - 1.0; }

template <int NT_, int ST_>
struct OpPlainCat : OpPlain<NT_, ST_> { static constexpr bool N_CAT = true; };   // 2 instructions per K step (engine header)

template <int NT, int ST, bool CAT = false>
static int run_plain(int M, int N, int K, int sm, bool timing) {
  std::vector<float> A((size_t)M * K), W((size_t)N * K);
  for (auto& x : A) x = (float)frand();
  for (auto& x : W) x = (float)(frand() * 0.5);
  const int npad = (N + NT - 1) / NT * NT;
  float *dA, *dW, *dWs, *dD;
  CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dW, W.size() * 4));
  CK(cudaMalloc(&dWs, (size_t)2 * npad * K * 4)); CK(cudaMalloc(&dD, (size_t)M * N * 4));
  CK(cudaMemset(dWs, 0, (size_t)2 * npad * K * 4));
  CK(cudaMemset(dD, 0xff, (size_t)M * N * 4));
  CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dW, W.data(), W.size() * 4, cudaMemcpyHostToDevice));
  split_kernel<<<(N * K + 255) / 256, 256>>>(dW, dWs, N, K, npad);
  using Op = typename std::conditional<CAT, OpPlainCat<NT, ST>, OpPlain<NT, ST>>::type;
  typename Op::Params p;
  if (!make_tmap_2d(&p.map, dWs, (uint64_t)2 * npad, (uint64_t)K, (uint64_t)K, NT)) { printf("tensor map failed\n"); return 1; }
  p.A = dA; p.D = dD; p.M = M; p.N = N; p.K = K; p.npad = npad; p.prof = nullptr; p.dbg = 0;
  p.mtiles = (M + MT - 1) / MT; p.ntiles_n = (N + NT - 1) / NT;
  CK(launch_gemm<Op>(p, p.mtiles * p.ntiles_n, sm, 0));
  CK(cudaDeviceSynchronize());
  std::vector<float> D((size_t)M * N);
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  double max_err = 0, max_ref = 0;
  for (int m = 0; m < M; m += (M > 512 ? 7 : 1))
    for (int n = 0; n < N; ++n) {
      double r = 0;
      for (int k = 0; k < K; ++k) r += (double)A[(size_t)m * K + k] * (double)W[(size_t)n * K + k];
      const double e = fabs((double)D[(size_t)m * N + n] - r);
      if (!(e <= max_err)) max_err = e;   // NaN-propagating
      if (fabs(r) > max_ref) max_ref = fabs(r);
    }
  printf("plain%s NT=%d ST=%d M=%d N=%d K=%d: max abs err %.3e / max |ref| %.3e = %.3e\n", CAT ? " N_CAT" : "", NT, ST, M, N, K, max_err, max_ref,
         max_err / max_ref);
  int bad = !(max_err / max_ref < 2e-6 + 2e-8 * K);
  if (timing) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    long long* dprof;
    CK(cudaMalloc(&dprof, sm * 16 * sizeof(long long)));
    for (int dbg = 0; dbg < 8; ++dbg) {
      p.dbg = dbg; p.prof = nullptr;
      for (int i = 0; i < 3; ++i) CK(launch_gemm<Op>(p, p.mtiles * p.ntiles_n, sm, 0));
      CK(cudaEventRecord(e0));
      const int reps = 10;
      for (int i = 0; i < reps; ++i) CK(launch_gemm<Op>(p, p.mtiles * p.ntiles_n, sm, 0));
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      const double fl = 2.0 * M * N * K * reps;
      CK(cudaMemset(dprof, 0, sm * 16 * sizeof(long long)));
      p.prof = dprof;
      CK(launch_gemm<Op>(p, p.mtiles * p.ntiles_n, sm, 0));
      CK(cudaDeviceSynchronize());
      std::vector<long long> pr(sm * 16);
      CK(cudaMemcpy(pr.data(), dprof, pr.size() * sizeof(long long), cudaMemcpyDeviceToHost));
      double w[5] = {0, 0, 0, 0, 0}, tot[4] = {0, 0, 0, 0};
      for (int b = 0; b < sm; ++b) {
        for (int r = 0; r < 5; ++r) w[r] += (double)pr[(b * 8 + r) * 2] / sm;
        for (int r = 0; r < 4; ++r) tot[r] += (double)pr[(b * 8 + r) * 2 + 1] / sm;
      }
      printf("  dbg=%d (skipA=%d skipTMA=%d skipStore=%d): %.3f ms, %.1f TFLOP/s fp32-eq | kclk per CTA: total %.0f; wait: tma %.0f mma-full %.0f mma-acc %.0f epi %.0f prod %.0f\n",
             dbg, dbg & 1, (dbg >> 1) & 1, (dbg >> 2) & 1, ms / reps, fl / (ms * 1e-3) / 1e12, tot[1] / 1e3, w[0] / 1e3, w[1] / 1e3,
             w[4] / 1e3, w[2] / 1e3, w[3] / 1e3);
    }
    p.dbg = 0; p.prof = nullptr;
    cudaFree(dprof);
  }
  cudaFree(dA); cudaFree(dW); cudaFree(dWs); cudaFree(dD);
  return bad;
}

template <int NT, int ST>
static int run_trans(int M, int N, int K, int sm) {
  std::vector<float> X((size_t)K * M), Y((size_t)K * N);
  for (auto& x : X) x = (float)frand();
  for (auto& x : Y) x = (float)frand();
  float *dX, *dY, *dD;
  CK(cudaMalloc(&dX, X.size() * 4)); CK(cudaMalloc(&dY, Y.size() * 4)); CK(cudaMalloc(&dD, (size_t)M * N * 4));
  CK(cudaMemset(dD, 0xff, (size_t)M * N * 4));
  CK(cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dY, Y.data(), Y.size() * 4, cudaMemcpyHostToDevice));
  using Op = OpTrans<NT, ST>;
  typename Op::Params p;
  p.X = dX; p.Y = dY; p.D = dD; p.M = M; p.N = N; p.K = K; p.prof = nullptr; p.dbg = 0;
  p.mtiles = (M + MT - 1) / MT; p.ntiles_n = (N + NT - 1) / NT;
  CK(launch_gemm<Op>(p, p.mtiles * p.ntiles_n, sm, 0));
  CK(cudaDeviceSynchronize());
  std::vector<float> D((size_t)M * N);
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  double max_err = 0, max_ref = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double r = 0;
      for (int k = 0; k < K; ++k) r += (double)X[(size_t)k * M + m] * (double)Y[(size_t)k * N + n];
      const double e = fabs((double)D[(size_t)m * N + n] - r);
      if (!(e <= max_err)) max_err = e;
      if (fabs(r) > max_ref) max_ref = fabs(r);
    }
  printf("trans NT=%d ST=%d M=%d N=%d K=%d: max abs err %.3e / max |ref| %.3e = %.3e\n", NT, ST, M, N, K, max_err, max_ref,
         max_err / max_ref);
  cudaFree(dX); cudaFree(dY); cudaFree(dD);
  return !(max_err / max_ref < 2e-6 + 2e-8 * K);
}

int main() {
  int dev = 0, sm = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
  printf("SMs: %d\n", sm);
  int bad = 0;
  bad += run_plain<256, 2>(128, 256, 32, sm, false);     // one tile, one chunk
  bad += run_plain<256, 2>(128, 256, 256, sm, false);    // one tile, 8 chunks
  bad += run_plain<256, 2>(384, 512, 256, sm, false);    // several tiles per CTA? (6 tiles on 148 SMs: one each)
  bad += run_plain<256, 2>(1000, 200, 72, sm, false);    // ragged M, N, K
  bad += run_plain<64, 4>(300, 40, 256, sm, false);      // small N tile
  bad += run_plain<256, 2>(128 * 600, 256, 256, sm, true);   // persistent loop: 600 tiles, timing
  bad += run_plain<128, 3>(128 * 600, 256, 256, sm, true);
  bad += run_plain<64, 4>(128 * 600, 64, 256, sm, true);
  bad += run_plain<64, 4>(128 * 600, 64, 64, sm, true);
  bad += run_plain<64, 4, true>(300, 40, 256, sm, false);
  bad += run_plain<64, 4, true>(1000, 64, 72, sm, false);
  bad += run_plain<64, 4, true>(128 * 600, 64, 256, sm, true);
  bad += run_plain<64, 4, true>(128 * 600, 64, 64, sm, true);
  bad += run_plain<64, 4, true>(128 * 1408, 64, 64, sm, true);
  bad += run_plain<64, 4>(128 * 1408, 64, 64, sm, true);
  bad += run_trans<128, 3>(256, 128, 96, sm);
  bad += run_trans<256, 2>(200, 250, 1000, sm);
  bad += run_trans<64, 4>(128, 64, 4096, sm);
  printf(bad ? "FAILED (%d)\n" : "ALL OK\n", bad);
  return bad ? 1 : 0;
}
