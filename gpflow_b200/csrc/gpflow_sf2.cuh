// gpflow_b200 -- "sf2": the 2-D stream-function train-step kernel (sm_100a), fp32, d = 2, nb_div_blocks = 1,
// two hidden layers of width <= 15.  This is the kernel behind BASELINE.json configs C1 / C2(no Euler) / C4.
//
// Same math and reference call sites as gpflow_small.cuh:
//   VectorField.eval/.jacobian  vector_field.py:75-128      divergence_free_vector_func  velocity_field.py:20-79
//   RK4 compute_step / ode_solve ode_order_method.py:66-71, ode_solver.py:70-81
//   Network.gp_flow network.py:156-189      Network.loss_gp_flow + backward  network.py:70-84,104-111
// For d = 2, nb = 1 the closed form A.2 only sees psi = f_0 - f_1 (stream function):
//   v_0 = s_0 (2 psi x_1 + s_1 d_1 psi),   v_1 = -s_1 (2 psi x_0 + s_0 d_0 psi),   s_j = x_j^2 - (1 - eps)
// so the output layer is collapsed to one row w3 = W3[0,:] - W3[1,:] when the weights are staged.
//
// Why a second kernel: the register-resident v1 kernel unrolls every contraction; its code (hundreds of KB) thrashes
// the 32 KB L1.5 instruction cache (ncu: no_instruction 4.8 stall cycles per issue, profiles/r1_v1_*).  Here
//   * one thread = one sample, one warp = a tile of 32 samples; contractions are ROLLED loops whose trip index is
//     the contraction index, accumulators stay in registers, per-sample activations live in shared-memory rows
//     [lane][feature] with an odd 16-byte stride (every access is a conflict-free 128-bit LDS/STS);
//   * the forward sweep does not recompute anything in the backward sweep: after each of the 4N velocity
//     evaluations the warp's activation rows (x, psi, grad psi, h1 | h2, U2) are streamed to an HBM stash by one
//     TMA bulk store (cp.async.bulk shared->global) and come back, double buffered, by TMA bulk loads + mbarrier
//     in the reverse sweep (9 KB per warp and evaluation, ~33 KB per sample and step, far below the HBM roof);
//   * the only cross-sample coupling, the weight gradients, is handled per layer: W2 (the 15x15 bulk) by a warp-
//     level rank-96 update per evaluation where each lane owns a 4x4 register tile of the gradient and reads the
//     other lanes' rows from shared memory; W1/b1 by a second, 16x4 rank update of the same kind; w3/b3 in registers;
//   * accumulation is two-level (per tile of 32 samples, then across tiles) so that 2^24-sample sums keep fp32
//     relative accuracy ~1e-6; the per-warp partial vectors are reduced in a fixed order by gpflow_finalize_kernel.
#pragma once
#include "gpflow_common.cuh"
#include "gpflow_small.cuh"  // gpflow_finalize_kernel, round_up

namespace gpflow {

struct SF2Args {
  const float* params;  // [P]
  const float* x_in;    // [B,2] x_map (wrap) or cube-space input
  const float* x_data;  // [B,2] OT target (train step)            } exactly one of the two is non-null
  const float* grad_y;  // [B,2] external cotangent (backward op)   }
  float* y_out;         // [B,2] nullable
  float* grad_x_out;    // [B,2] nullable
  float* partials;      // [nslots][npart]: [3 scalars | P grads]
  float* g2;            // [nslots][NACC][32] second-level accumulators
  char* stash;          // [nslots][4N][STASH_BYTES]
  long long B;
  int n_steps;
  int npart;
  unsigned flags;
  float dt;  // signed solve dt
  float t0;
  float sigma;
  float loss_scale;
  int stagger_cycles;  // sf2r: warps of the upper half start this many cycles late (GPFLOW_SF2R_STAGGER, experiments)
  int cta_align;        // sf2e: 1 = named barrier over the CTA's active warps before every RK4 step (keeps them in phase)
};

// weights never change after the prologue: volatile asm keeps each (warp-uniform) load next to its use instead of
// letting the compiler hoist hundreds of them out of the time loop
GPF_DEV float4 ldsw4(const float* p) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(smem_u32(p)));
  return v;
}
GPF_DEV float ldsw1(const float* p) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(smem_u32(p)));
  return v;
}
// GPF_SF2_CONST: 0 = packed weights in shared memory (LDS.128, 2 shared-memory wavefronts per warp-uniform load),
// 1 = packed weights in the constant bank (no shared-memory traffic; with unrolled contractions the FFMA reads its
// weight operand straight from c[][] and the load instruction disappears).
#ifndef GPF_SF2_CONST
#define GPF_SF2_CONST 0
#endif
#ifndef GPF_SF2_UNROLL
#define GPF_SF2_UNROLL 0
#endif
#if GPF_SF2_CONST
__constant__ float sf2_cw[640];
GPF_DEV float4 ldw4(const float*, int off) { return *reinterpret_cast<const float4*>(&sf2_cw[off]); }
#else
GPF_DEV float4 ldw4(const float* W, int off) { return ldsw4(W + off); }
#endif
#if GPF_SF2_UNROLL
#define GPF_SF2_LOOP _Pragma("unroll")
#else
#define GPF_SF2_LOOP _Pragma("unroll 1")
#endif
GPF_DEV float4 lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }
GPF_DEV void sts4(float* p, float a, float b, float c, float d) { *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d); }

GPF_DEV void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}
GPF_DEV void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> GPF_DEV void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> GPF_DEV void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
GPF_DEV void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

template <int H1_, int H2_, int NW_>
struct SF2 {
  static constexpr int H1 = H1_, H2 = H2_, NW = NW_;
  static_assert(H1 >= 1 && H1 <= 15 && H2 >= 1 && H2 <= 15, "slot 15 of the 16-wide rows is reserved");
  // flat parameter layout (vector_field.py:56-61), d = 2, nb = 1
  static constexpr int OFF_W1 = 0, OFF_B1 = 3 * H1, OFF_W2 = OFF_B1 + H1, OFF_B2 = OFF_W2 + H2 * H1,
                       OFF_W3 = OFF_B2 + H2, OFF_B3 = OFF_W3 + 2 * H2, NPARAM = OFF_B3 + 2;
  // packed weights (floats)
  static constexpr int W1P = 0;           // [16][4]  {W1[k][0], W1[k][1], W1[k][2] (time), b1[k]}; rows >= H1 zero
  static constexpr int W2T = W1P + 64;    // [16][16] W2T[m][k] = W2[k][m]; row m = 15 holds b2 (h1 slot 15 == 1)
  static constexpr int W2R = W2T + 256;   // [16][16] W2R[k][m] = W2[k][m]
  static constexpr int W3C = W2R + 256;   // [16]     W3[0][k] - W3[1][k]
  static constexpr int B3C = W3C + 16;    // [4]      b3[0] - b3[1], 0, 0, 0
  static constexpr int WEND = B3C + 4;
  // per-warp rows
  static constexpr int AS = 20;  // A row: x0 x1 psi g0 | h1[0..14] 1.0          (80 B  = 5 x 16)
  static constexpr int BS = 52;  // B row: h2[0..14] g1 | U2_0[16] | U2_1[16] | pad (208 B = 13 x 16)
  static constexpr int A_FL = 32 * AS, B_FL = 32 * BS;
  static constexpr int A_BYTES = A_FL * 4, B_BYTES = B_FL * 4;
  static constexpr int STASH_BYTES = A_BYTES + B_BYTES;  // per warp and evaluation
  static constexpr int WARP_FL = 2 * A_FL + 2 * B_FL;
  static constexpr int NACC = 50;  // second level: 16 W2 tile | 16 w3 | b3 | 16 W1 tile | ot
  static constexpr int G_W2 = 0, G_W3 = 16, G_B3 = 32, G_W1 = 33, G_OT = 49;
  static constexpr int SCT = 16;   // stage-constant table [4][4] {t offset, k weight, y step, k weight * dt/6}

  static size_t smem_bytes(int n_steps) {
    return (size_t)(WEND + SCT + round_up(n_steps + 1, 4) + round_up(NPARAM, 4) + NW * WARP_FL) * sizeof(float) + 128;
  }
};

}  // namespace gpflow
