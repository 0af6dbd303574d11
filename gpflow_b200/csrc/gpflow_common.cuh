// gpflow_b200 -- shared device helpers (sm_100a).
//
// Scalar algebra used by every kernel: the velocity evaluation and its reverse sweep are written
// once, templated on a scalar type S that is either the real type T (float / double) or Dual<T>
// (value + one epsilon part).  Dual is what turns the Euler penalization (res_net.py:319-358) into
// the same code: J_v w is the epsilon part of v(x + eps w), and the gradient of a functional of
// (v, J_v w) is the epsilon part of the SAME reverse sweep run in dual arithmetic with the adjoint
// seed (adjoint-of-tangent) + eps (adjoint-of-value)  [reverse-over-forward; DESIGN.md "Euler"].
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gpflow {

#define GPF_DEV __device__ __forceinline__

template <typename T>
struct Dual {
  T v, e;
  GPF_DEV Dual() {}
  GPF_DEV Dual(T v_) : v(v_), e(T(0)) {}
  GPF_DEV Dual(T v_, T e_) : v(v_), e(e_) {}
};

template <typename S> struct RealOf { typedef S type; };
template <typename T> struct RealOf<Dual<T> > { typedef T type; };
template <typename S> struct IsDual { static constexpr bool value = false; };
template <typename T> struct IsDual<Dual<T> > { static constexpr bool value = true; };

template <typename T> GPF_DEV Dual<T> operator+(Dual<T> a, Dual<T> b) { return Dual<T>(a.v + b.v, a.e + b.e); }
template <typename T> GPF_DEV Dual<T> operator-(Dual<T> a, Dual<T> b) { return Dual<T>(a.v - b.v, a.e - b.e); }
template <typename T> GPF_DEV Dual<T> operator-(Dual<T> a) { return Dual<T>(-a.v, -a.e); }
template <typename T> GPF_DEV Dual<T> operator*(Dual<T> a, Dual<T> b) { return Dual<T>(a.v * b.v, a.v * b.e + a.e * b.v); }
template <typename T> GPF_DEV Dual<T> operator*(T a, Dual<T> b) { return Dual<T>(a * b.v, a * b.e); }
template <typename T> GPF_DEV Dual<T> operator*(Dual<T> a, T b) { return Dual<T>(a.v * b, a.e * b); }
template <typename T> GPF_DEV Dual<T> operator+(T a, Dual<T> b) { return Dual<T>(a + b.v, b.e); }
template <typename T> GPF_DEV Dual<T> operator+(Dual<T> a, T b) { return Dual<T>(a.v + b, a.e); }
template <typename T> GPF_DEV Dual<T> operator-(T a, Dual<T> b) { return Dual<T>(a - b.v, -b.e); }
template <typename T> GPF_DEV Dual<T> operator-(Dual<T> a, T b) { return Dual<T>(a.v - b, a.e); }
template <typename T> GPF_DEV Dual<T>& operator+=(Dual<T>& a, Dual<T> b) { a.v += b.v; a.e += b.e; return a; }
template <typename T> GPF_DEV Dual<T>& operator-=(Dual<T>& a, Dual<T> b) { a.v -= b.v; a.e -= b.e; return a; }

// component access: comp<0> = value part, comp<1> = epsilon part (reals have no epsilon part)
template <int C, typename T> GPF_DEV T comp(const Dual<T>& a) { return C == 0 ? a.v : a.e; }
template <int C> GPF_DEV float comp(const float& a) { return a; }
template <int C> GPF_DEV double comp(const double& a) { return a; }

// ---- elementary functions (accurate variants: no tanh.approx / fast-math) ----
// tanh, fp32.  GPF_TANH_IMPL: 0 = CUDA libm tanhf (branchy), 1 = branch-free two-regime evaluation with libm-class
// accuracy (default): |x| < 0.55 -> x + x^3 P(x^2), P degree 4 fitted on [0, 0.6] (max error 0.61 ulp in fp32
// arithmetic); otherwise 1 - 2/(2^(2x log2 e) + 1) on MUFU.EX2 / MUFU.RCP (<= ~2 ulp).  No tanh.approx (2^-11).
#ifndef GPF_TANH_IMPL
#define GPF_TANH_IMPL 1
#endif
GPF_DEV float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
GPF_DEV float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
GPF_DEV float tanh_real(float x) {
#if GPF_TANH_IMPL == 0
  return tanhf(x);
#else
  const float ax = fabsf(x);
  const float u = x * x;
  float p = fmaf(u, -0.005919910501688719f, 0.02081829123198986f);
  p = fmaf(p, u, -0.05379083380103111f);
  p = fmaf(p, u, 0.13332012295722961f);
  p = fmaf(p, u, -0.33333301544189453f);
  const float small = fmaf(x * u, p, x);
  const float e = ex2_approx(ax * 2.885390043258667f);
  const float r = rcp_approx(e + 1.0f);
  const float big = copysignf(fmaf(-2.0f, r, 1.0f), x);
  return ax < 0.55f ? small : big;
#endif
}
GPF_DEV double tanh_real(double x) { return tanh(x); }
// Two tanh_real at once with the polynomial regime, the `e + 1` and the `1 - 2 r` of the exponential regime as packed
// fp32x2 instructions (sm_100 FFMA2 / FMUL2 / FADD2): the SAME fp32 operations per element as tanh_real, hence bit-identical
// results, in 21 instead of 32 issue slots (the d = 2 kernels are issue / FMA-pipe bound, profiles/r2_sf2r_ncu_summary.txt).
GPF_DEV float2 tanh_real2(float2 x) {
#if GPF_TANH_IMPL == 0
  return make_float2(tanhf(x.x), tanhf(x.y));
#else
  const float2 u = __fmul2_rn(x, x);
  float2 p = __ffma2_rn(u, make_float2(-0.005919910501688719f, -0.005919910501688719f),
                        make_float2(0.02081829123198986f, 0.02081829123198986f));
  p = __ffma2_rn(p, u, make_float2(-0.05379083380103111f, -0.05379083380103111f));
  p = __ffma2_rn(p, u, make_float2(0.13332012295722961f, 0.13332012295722961f));
  p = __ffma2_rn(p, u, make_float2(-0.33333301544189453f, -0.33333301544189453f));
  const float2 small = __ffma2_rn(__fmul2_rn(x, u), p, x);
  const float ax = fabsf(x.x), ay = fabsf(x.y);
  const float2 e = make_float2(ex2_approx(ax * 2.885390043258667f), ex2_approx(ay * 2.885390043258667f));
  const float2 e1 = __fadd2_rn(e, make_float2(1.0f, 1.0f));
  const float2 r = make_float2(rcp_approx(e1.x), rcp_approx(e1.y));
  const float2 big = __ffma2_rn(make_float2(-2.0f, -2.0f), r, make_float2(1.0f, 1.0f));
  return make_float2(ax < 0.55f ? small.x : copysignf(big.x, x.x), ay < 0.55f ? small.y : copysignf(big.y, x.y));
#endif
}
// tanh as 1 - 2 / (2^(2 x log2 e) + 1) on MUFU.EX2 / MUFU.RCP for every x (no polynomial branch, no sign handling:
// the formula saturates to -1 / +1 by itself): 3 FMA-pipe instructions instead of 10.  ABSOLUTE error <= ~3e-7
// (2-3 ulp of 1.0; ex2.approx 2^-22 relative on a value near 1, rcp.approx 1 ulp), i.e. the relative accuracy of tanh_real
// is given up for |x| << 1.  NOT used by default: at the reference's initialisation (weights ~1e-3) every argument is
// tiny and the training-trajectory parity tests fail with it (see GPF_SF2R_TANH in gpflow_sf2r.cuh); kept for A/B runs.
GPF_DEV float tanh_exp2(float x) {
  const float e = ex2_approx(x * 2.885390043258667f);
  return fmaf(-2.0f, rcp_approx(e + 1.0f), 1.0f);
}
GPF_DEV float erf_real(float x) { return erff(x); }
GPF_DEV double erf_real(double x) { return erf(x); }
GPF_DEV float erfinv_real(float x) { return erfinvf(x); }
GPF_DEV double erfinv_real(double x) { return erfinv(x); }
GPF_DEV float exp_real(float x) { return expf(x); }
GPF_DEV double exp_real(double x) { return exp(x); }

GPF_DEV float act_tanh(float a) { return tanh_real(a); }
GPF_DEV double act_tanh(double a) { return tanh_real(a); }
template <typename T> GPF_DEV Dual<T> act_tanh(Dual<T> a) {
  T h = tanh_real(a.v);
  return Dual<T>(h, (T(1) - h * h) * a.e);
}

template <typename T> struct Eps;
template <> struct Eps<float> { static constexpr float value = 1.1920928955078125e-07f; };   // np.finfo(float32).eps
template <> struct Eps<double> { static constexpr double value = 2.220446049250313e-16; };   // np.finfo(float64).eps

template <typename T> GPF_DEV T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <typename T> GPF_DEV Dual<T> warp_sum(Dual<T> a) { return Dual<T>(warp_sum(a.v), warp_sum(a.e)); }
template <typename T> GPF_DEV T shfl_xor_s(T v, int o) { return __shfl_xor_sync(0xffffffffu, v, o); }
template <typename T> GPF_DEV Dual<T> shfl_xor_s(Dual<T> a, int o) {
  return Dual<T>(__shfl_xor_sync(0xffffffffu, a.v, o), __shfl_xor_sync(0xffffffffu, a.e, o));
}
// Four warp sums in 6 shuffles instead of 20: the first two butterfly stages exchange half of the values, so that
// lane L ends with the warp total of v[2 * ((L >> 4) & 1) + ((L >> 3) & 1)].
template <typename S> GPF_DEV S warp_sum4(const S (&v)[4], int lane) {
  const bool h16 = (lane & 16) != 0, h8 = (lane & 8) != 0;
  S k0 = h16 ? v[2] : v[0], k1 = h16 ? v[3] : v[1];
  const S s0 = h16 ? v[0] : v[2], s1 = h16 ? v[1] : v[3];
  k0 = k0 + shfl_xor_s(s0, 16);
  k1 = k1 + shfl_xor_s(s1, 16);
  S k = h8 ? k1 : k0;
  const S s = h8 ? k0 : k1;
  k = k + shfl_xor_s(s, 8);
  k = k + shfl_xor_s(k, 4);
  k = k + shfl_xor_s(k, 2);
  k = k + shfl_xor_s(k, 1);
  return k;
}

__host__ __device__ constexpr int round_up(int a, int b) { return (a + b - 1) / b * b; }
__host__ __device__ constexpr int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---- TMA bulk copy (cp.async.bulk, SASS UBLKCP) + mbarrier: stage the flat parameters once per CTA ----
GPF_DEV uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
GPF_DEV void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
GPF_DEV void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
GPF_DEV void mbar_wait(uint64_t* bar, uint32_t phase) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(phase)
        : "memory");
  }
}
GPF_DEV void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
GPF_DEV void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
GPF_DEV void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Copy n elements of T from global to shared, TMA bulk for the 16-byte aligned body, plain loads for the tail.
// Must be called by all threads of the CTA; `bar` is a shared mbarrier initialised by this routine.
template <typename T>
__device__ inline void stage_params_to_smem(T* dst, const T* src, int n, uint64_t* bar) {
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
  const uint32_t body_bytes = aligned ? (uint32_t)((size_t)n * sizeof(T)) & ~15u : 0u;
  const int body = (int)(body_bytes / sizeof(T));
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0 && body_bytes > 0) {
    mbar_expect_tx(bar, body_bytes);
    // bulk copies are limited in size per instruction only by the tx-count width (2^20-1 bytes); chunk anyway
    uint32_t off = 0;
    while (off < body_bytes) {
      uint32_t chunk = body_bytes - off;
      if (chunk > 65536u) chunk = 65536u;
      bulk_g2s(reinterpret_cast<char*>(dst) + off, reinterpret_cast<const char*>(src) + off, chunk, bar);
      off += chunk;
    }
  }
  for (int i = body + threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  if (body_bytes > 0) mbar_wait(bar, 0);
  __syncthreads();
}

}  // namespace gpflow
