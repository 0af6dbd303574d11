/*
 * gpflow_b200 -- C ABI of the B200-native GP-flow (Gaussian-preserving flow) hot path.
 *
 * The reference (morel-g/GPFlow) is pure Python/PyTorch and has no FFI layer; the boundary this
 * library sits behind is the nn.Module contract of src/gp_flows/res_net.py.  Each
 * entry point below names the reference call sites whose work it replaces.  The host side
 * (gpflow_b200/res_net.py, gpflow_b200/network.py) mirrors those modules and reaches this ABI
 * through ctypes with raw device pointers; see INTEGRATION.md for the binding.
 *
 * Contract
 *   - plain C: pointers, sizes, one POD descriptor.  No torch / C++ types cross the boundary.
 *   - every pointer is a DEVICE pointer of the current CUDA device unless said otherwise; all
 *     arrays are dense row-major of the element type selected by GpFlowDesc.dtype.
 *   - the caller owns every buffer including the workspace; the library never allocates or frees
 *     device memory and keeps no reference to caller memory after the call returns.
 *   - all work is enqueued on the given stream; no host<->device synchronisation, no host read of
 *     device data => calls are CUDA-graph capturable.  Re-entrant (no mutable global state).
 *   - return value 0 = OK, otherwise a GpFlowStatus; no C++ exception crosses the boundary.
 *     NaNs are data, not errors (the reference's NaN guard lives in validation_step,
 *     src/gp_flows/network.py:134-138).
 */
#ifndef GPFLOW_B200_H
#define GPFLOW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPFLOW_ABI_VERSION 1
#define GPFLOW_MAX_HIDDEN 8

typedef enum GpFlowStatus {
  GPFLOW_OK = 0,
  GPFLOW_ERR_BAD_DESC = 1,       /* inconsistent / out-of-range descriptor field            */
  GPFLOW_ERR_UNSUPPORTED = 2,    /* valid but outside what the kernels implement             */
  GPFLOW_ERR_NULL_POINTER = 3,   /* a required pointer is NULL                               */
  GPFLOW_ERR_WORKSPACE = 4,      /* workspace smaller than gpflow_query_workspace() reports  */
  GPFLOW_ERR_MISALIGNED = 5,     /* a buffer is not aligned to its element size              */
  GPFLOW_ERR_CUDA = 6,           /* cudaGetLastError() after a launch was not cudaSuccess    */
  GPFLOW_ERR_STATIC_SCRATCH = 7, /* gpflow_velocity_static(): the call needs more scratch than
                                    GPFLOW_STATIC_SCRATCH_BYTES; use gpflow_velocity() with a workspace */
  GPFLOW_ERR_COLLECTIVE = 8      /* gpflow_comm_* / gpflow_allreduce_grads: NCCL not found at run time, or it returned an error */
} GpFlowStatus;

typedef enum GpFlowDType { GPFLOW_F32 = 0, GPFLOW_F64 = 1 } GpFlowDType;

typedef enum GpFlowOp { GPFLOW_OP_FORWARD = 0, GPFLOW_OP_TRAIN_STEP = 1, GPFLOW_OP_BACKWARD = 2 } GpFlowOp;

/* GpFlowDesc.flags */
#define GPFLOW_FLAG_WRAP_GAUSS 0x1u /* input lives in Gaussian space: g0 = erf(x/(sqrt2 sigma)) in front, y = sqrt2 sigma
                                       erfinv(clip(g_T, +-(1-eps))) behind  (network.py:174-182).  Without it the op is the
                                       raw cube-space flow, i.e. ResNet.forward (res_net.py:178-197).               */
#define GPFLOW_FLAG_LAGRANGIAN 0x2u /* accumulate dt * mean_b sum_i 1/2 v_i(x_{k+1}, t_k)^2 (res_net.py:292-294,360-361) */
#define GPFLOW_FLAG_EULER_PEN 0x4u  /* Euler penalization (res_net.py:295-302,319-358); needs zw probes              */
#define GPFLOW_FLAG_REVERSE 0x8u    /* integrate from t0 = dt*n_steps_nominal with -dt (res_net.py:217-219)          */
#define GPFLOW_FLAG_FORCE_GENERIC 0x10u /* testing aid: bypass the specialised small-d kernels                         */
#define GPFLOW_FLAG_FORCE_V1 0x20u      /* testing aid: use the register-resident v1 small-d kernel instead of sf2     */
#define GPFLOW_FLAG_FORCE_TC 0x40u      /* testing aid: use the chunked tensor-core (tcgen05) path whatever the batch size;
                                           GPFLOW_ERR_UNSUPPORTED if the shape is outside it (fp32, no Euler, d <= 64,
                                           hidden widths <= 256).  Without it the path is taken above a measured
                                           batch threshold (128 for wide nets, 6144 for widths <= 64).           */

/*
 * One GP flow = get_gp_flow(dim, velocity_dict, nb_layers_gp, T_final, ...) with nb_blocks == 1
 * (res_net.py:14-67).  Only what the kernels implement is representable: tanh activation,
 * continuous-time, incompressible velocity, RK4 / Case.standard; everything else stays on the
 * reference's Python classes (SURVEY.md 8b "Unsupported by the kernels").
 */
typedef struct GpFlowDesc {
  int32_t abi_version;            /* GPFLOW_ABI_VERSION                                                          */
  int32_t dtype;                  /* GpFlowDType: precision mode (src/precision.py:7-13); also fixes eps         */
  int32_t d;                      /* data.dim                                                                    */
  int32_t nb;                     /* velocity_dict["nb_div_blocks"]  (1 <= nb <= d)                              */
  int32_t n_hidden;               /* len(velocity_dict["nb_neurons"]) (1..GPFLOW_MAX_HIDDEN)                     */
  int32_t hidden[GPFLOW_MAX_HIDDEN]; /* velocity_dict["nb_neurons"]                                              */
  int32_t n_steps;                /* RK4 steps integrated: nb_layers (training) or nb_layers_eval (res_net.py:209-214) */
  int32_t n_steps_nominal;        /* self.nb_layers; only used for the reverse start time (res_net.py:218)        */
  uint32_t flags;                 /* GPFLOW_FLAG_*                                                               */
  int32_t reserved0;
  double dt;                      /* dt_solve > 0: self.dt or self.dt_eval                                       */
  double dt_constraint;           /* self.dt -- the Lagrangian's weight is always self.dt (res_net.py:293)       */
  double sigma;                   /* Network.sigma (network.py:51), used with GPFLOW_FLAG_WRAP_GAUSS             */
  double euler_coef;              /* flow.euler_coef_penalization for this call (mutated by GPFlowCallback)      */
  double loss_scale;              /* 1 / B_global.  Every sum over samples this library returns (OT cost, penalty,
                                     Lagrangian, parameter gradients of the fused train step) is multiplied by it, so
                                     that a SUM over data-parallel ranks gives the reference's batch MEAN.       */
} GpFlowDesc;

/* Number of parameters: sum_l (n_l * n_{l+1} + n_{l+1}), n = [d+1, hidden..., d*nb].  The flat parameter /
 * gradient layout is [W_1, b_1, ..., W_{L+1}, b_{L+1}], each W row-major [out, in] exactly as
 * nn.Linear.weight (vector_field.py:56-61), i.e. torch.cat([p.reshape(-1) for p in flow.parameters()]). */
int64_t gpflow_param_count(const GpFlowDesc* desc);

/* scalars_out layout (3 elements of dtype): */
#define GPFLOW_SCALAR_OT 0   /* loss_scale * sum_b ||x_data_b - y_b||^2        (network.py:82)      */
#define GPFLOW_SCALAR_REG 1  /* sum_k euler_coef * loss_scale * sum_b q_bk^2   (res_net.py:300-302)  */
#define GPFLOW_SCALAR_LAGR 2 /* dt_constraint * loss_scale * sum_k sum_b 1/2|v|^2 (res_net.py:292-294) */
#define GPFLOW_NUM_SCALARS 3

/* Bytes of caller-provided workspace `op` needs for a batch of B samples (16-byte aligned start). */
int gpflow_query_workspace(const GpFlowDesc* desc, int64_t B, int op, size_t* bytes);

/*
 * Forward transport.  Replaces ResNet.forward/solve/process_ode_solution (res_net.py:178-272) and, with
 * GPFLOW_FLAG_WRAP_GAUSS, Network.gp_flow (network.py:156-189).
 *   params      [P]        flat parameters
 *   x_in        [B,d]      cube-space points (raw) or Gaussian-space points (WRAP_GAUSS)
 *   zw          [n_steps,2,d] Euler probes z,w per step (nullable unless EULER_PEN)
 *   y_out       [B,d]      transported points
 *   logdet_out  [B]        nullable; (||y||^2 - ||x||^2) / (2 sigma^2), WRAP_GAUSS only (network.py:183-185)
 *   traj_out    [n_steps+1,B,d] nullable; every RK4 state z_k in cube space (ode_solver.py:70-81)
 *   scalars_out [3]        nullable; see GPFLOW_SCALAR_* (OT slot is written as 0)
 */
int gpflow_forward(const GpFlowDesc* desc, int64_t B, const void* params, const void* x_in, const void* zw, void* y_out,
                   void* logdet_out, void* traj_out, void* scalars_out, void* workspace, size_t workspace_bytes,
                   void* cuda_stream);

/*
 * Fused train step: forward, OT cost, Euler penalty, and the full backward pass in one launch (+ a
 * deterministic finalize).  Replaces Network.loss_gp_flow + regularization_coef + manual_backward of the GP
 * branch of training_step (network.py:70-84,104-111) for everything after the frozen map.
 *   x_data  [B,d]   data points x (OT target)          x_map [B,d]  frozen-map output map(x)[0]
 *   grads_out [P]   d(loss_scale*sum_b ||x-y||^2 + reg)/d params, flat layout above
 *   scalars_out [3] GPFLOW_SCALAR_*
 *   y_out [B,d] nullable.   grad_xmap_out [B,d] nullable: d loss / d x_map (ot_map(no_grad_map=False), network.py:227-236)
 * Requires GPFLOW_FLAG_WRAP_GAUSS.
 */
int gpflow_train_step(const GpFlowDesc* desc, int64_t B, const void* params, const void* x_data, const void* x_map,
                      const void* zw, void* grads_out, void* scalars_out, void* y_out, void* grad_xmap_out,
                      void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * Generic backward for the autograd case (any loss on top of the flow): given the cotangent of the
 * transported points, recompute the trajectory and run the reverse sweep.
 *   grad_y       [B,d]  cotangent of y_out (already carrying whatever batch normalisation the loss has)
 *   reg_cotangent       nullable device scalar: cotangent of regularization_value (NULL = 1); the Euler term
 *                       contributes reg_cotangent * euler_coef * loss_scale * sum_b q^2 per step
 *   grads_out [P], grad_x_out [B,d] (nullable)
 */
int gpflow_backward(const GpFlowDesc* desc, int64_t B, const void* params, const void* x_in, const void* grad_y,
                    const void* reg_cotangent, const void* zw, void* grads_out, void* grad_x_out,
                    void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * Velocity field evaluation v(x, t): VelocityField.__call__ (velocity_field.py:118-126), used outside the train step
 * (trajectory / quiver plots, res_net.py:248-272).  x, v_out [B,d].  Workspace: gpflow_query_workspace(.., OP_FORWARD)
 * with GPFLOW_FLAG_FORCE_GENERIC set in the descriptor (the evaluation always runs on the generic kernel family).
 */
int gpflow_velocity(const GpFlowDesc* desc, int64_t B, const void* params, const void* x, double t, void* v_out,
                    void* workspace, size_t workspace_bytes, void* cuda_stream);

/*
 * Standalone VectorField.eval / VectorField.jacobian (src/gp_flows/vector_field.py:75-128; reference test
 * tests/test_vector_field.py:15-25): the tanh MLP with `n_linear` nn.Linear layers of widths
 * widths[0] (= dim_in, the velocity field passes [x; t]) ... widths[n_linear] (= dim_out), flat parameters
 * [W_1, b_1, ...] as everywhere else.  f_out [B, dim_out] and / or jac_out [B, dim_out, dim_in] (either may be NULL).
 * Not on the hot path: the fused kernels evaluate both inside the velocity.  One launch, no workspace;
 * GPFLOW_ERR_UNSUPPORTED when two layers of activations + tangents exceed 200 KB of shared memory.
 */
int gpflow_potentials(int dtype, int n_linear, const int32_t* widths, int64_t B, const void* params, const void* x,
                      void* f_out, void* jac_out, void* cuda_stream);

/*
 * gpflow_velocity() for callers that cannot allocate or pass a workspace (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as gpflow_velocity() on the generic kernel family; the scratch is a
 * file-scope device array of GPFLOW_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds
 * it until the process exits).  The call allocates nothing: no cudaMalloc / cudaFree, no allocating library routine.
 *   - limit: GPFLOW_ERR_STATIC_SCRATCH when gpflow_query_workspace(desc | GPFLOW_FLAG_FORCE_GENERIC, B, OP_FORWARD)
 *     exceeds GPFLOW_STATIC_SCRATCH_BYTES (d=2 H=[15,15] and d=5 H=[5,5,5]: any B; d=10 nb=9 H=[50,50,50] fp32:
 *     B <= 256); nothing is launched in that case;
 *   - every call shares the one array: calls that OVERLAP on the GPU (other streams, other threads) race on it; calls
 *     enqueued on one stream are safe.  It keeps the previous call's contents; the evaluation does not depend on them;
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress, may load the module):
 *     make one call per device before starting a graph capture.
 */
#define GPFLOW_STATIC_SCRATCH_BYTES (32u << 20)
int gpflow_velocity_static(const GpFlowDesc* desc, int64_t B, const void* params, const void* x, double t, void* v_out,
                           void* cuda_stream);

/*
 * Multi-GPU hook (SURVEY.md 8b "Multi-GPU hook", 8e): the path shards by samples -- every rank runs gpflow_train_step on its
 * contiguous batch shard with desc->loss_scale = 1 / B_global, which leaves SUMS already scaled by 1 / B_global in the fused
 * buffer [grads (P) | ot, reg, lagr]; ONE in-place SUM all-reduce of those P + 3 elements yields the global-batch loss and
 * gradient on every rank (gpflow_b200/data_parallel.py does the same through torch.distributed).  The reference is
 * single-device (training.py:124-136), so there is no reference interface to replace; these four entry points let a host
 * without torch.distributed do the exchange.  NCCL is bound at run time (dlopen "libnccl.so.2", or $GPFLOW_NCCL_LIB; an
 * already loaded copy -- PyTorch's -- is preferred): the library itself links against the CUDA runtime only.
 *   gpflow_comm_unique_id   rank 0: fills 128 bytes that the host distributes to the other ranks (any transport);
 *   gpflow_comm_create      every rank, on its own device (cudaSetDevice first): collective; returns an opaque communicator;
 *   gpflow_allreduce_grads  in-place SUM over the ranks of buf[0..count) (dtype GPFLOW_F32 / GPFLOW_F64), enqueued on
 *                           `cuda_stream` behind the train step that wrote buf; no host synchronisation;
 *   gpflow_comm_destroy     releases the communicator.
 * GPFLOW_ERR_COLLECTIVE when NCCL cannot be loaded or reports an error.
 */
int gpflow_comm_unique_id(void* id_out_128_bytes);
int gpflow_comm_create(const void* id_128_bytes, int n_ranks, int rank, void** comm_out);
int gpflow_allreduce_grads(void* buf, int64_t count, int dtype, void* comm, void* cuda_stream);
int gpflow_comm_destroy(void* comm);

/* Name of the kernel family an op on this descriptor dispatches to ("small<f32,2,1,15,15>", "generic<f64>", ...). */
int gpflow_kernel_variant(const GpFlowDesc* desc, int op, char* buf, size_t buf_len);

/* Number of kernel launches one call of `op` enqueues (for launch accounting in benchmarks). */
int gpflow_launch_count(const GpFlowDesc* desc, int op);

/* Launch accounting of the chunked tensor-core path (whose launch count depends on the batch: slabs x evaluations x
 * kernels per evaluation): cumulative number of kernels the path has enqueued in this process (host counter; take the
 * difference around the region of interest).  gpflow_tc_slab_samples: samples per slab for a batch of B (-1 if the
 * shape is outside the path). */
int64_t gpflow_tc_launches_total(void);
int gpflow_tc_slab_samples(const GpFlowDesc* desc, int64_t B, int op);

/* FP32 FMA micro-benchmark (roofline denominator): every thread of a grid that fills all SMs runs
 * `iters` x 128 independent-chain FFMAs.  `sink` is a device float[1] (never written in practice).  The number of
 * flops enqueued (2 per FFMA) is returned through `host_flops_out` (HOST pointer).  Time it with CUDA events. */
int gpflow_fp32_fma_burn(int64_t iters, void* sink, double* host_flops_out, void* cuda_stream);

/* TF32 tensor-pipe micro-benchmark: every warp of a grid that fills all SMs issues `iters` x 8 independent
 * mma.sync.m16n8k8 tf32 (fp32 accumulate), the instruction the generic kernel's 3xTF32 contractions use.  The dense
 * tf32 flops enqueued are returned through `host_flops_out` (HOST pointer); the fp32-equivalent roof of a 3xTF32
 * contraction is one third of the measured rate. */
int gpflow_tf32_mma_burn(int64_t iters, void* sink, double* host_flops_out, void* cuda_stream);

/* SM clock probe for benchmarks: one thread spins ~40k SM cycles and writes {clock64 ticks, globaltimer ns} to the
 * device uint64[2] `out2`; SM MHz = 1e3 * ticks / ns.  Enqueued between timed steps it samples the clock the GPU is
 * actually running at under load without the driver round trip of an NVML / nvidia-smi query. */
int gpflow_clock_probe(void* out2, void* cuda_stream);

/* Fused Adam step on the flat parameter buffer -- the `gp_flow_opt.step()` of the GP branch of
 * `Network.training_step` (src/gp_flows/network.py:104-111) for the optimiser built in `configure_optimizers`
 * (network.py:307-317: `torch.optim.Adam(flow_params, lr, betas=(0.9, 0.999), weight_decay)`, L2-coupled decay, no
 * amsgrad).  Same update as torch.optim.Adam:
 *     g = grad + weight_decay * p;  m = m + (1 - beta1) (g - m);  v = beta2 v + (1 - beta2) g^2
 *     p = p - lr / (1 - beta1^step) * m / (sqrt(v) / sqrt(1 - beta2^step) + eps)
 * `step` is the 1-based count of this update.  The StepLR schedule (network.py:318-320) is a host-side scalar: pass
 * the current `lr`.  dtype: GPFLOW_F32 / GPFLOW_F64; all buffers have `n` elements; `params`, `exp_avg`, `exp_avg_sq`
 * are updated in place.  One launch. */
int gpflow_adam_step(int dtype, int64_t n, void* params, const void* grads, void* exp_avg, void* exp_avg_sq, double lr,
                     double beta1, double beta2, double eps, double weight_decay, int64_t step, void* cuda_stream);

/* Forward of the frozen BNAF map in front of the GP flow: `x_map, log_det = self.map(x)` of `Network.loss_gp_flow`
 * (src/gp_flows/network.py:76-78) for `flow_models/bnaf/bnaf.py:11-39` with ONE block (`nblocks = 1`, res = None, the
 * shape of pretrained_models/{eight_gaussians,moons}/bnaf): MaskedWeight(dim -> H) Tanh [MaskedWeight(H -> H) Tanh]^(n-2)
 * MaskedWeight(H -> dim), H = dim * hidden_per_dim, n = n_layers masked linear layers (bnaf_module.py:151-325).
 * `params` = concatenation over the masked layers, in state_dict order, of `_weight [out][in]`, `_diag_weight [out]`,
 * `bias [out]` (the `mask_d` / `mask_o` buffers are implied by the shape).  fp32 only; supported shapes: dim = 2,
 * hidden_per_dim = 16; anything else returns GPFLOW_ERR_UNSUPPORTED (the caller keeps the reference module).
 * Writes y_out [B, dim] and, if not NULL, logdet_out [B].  One launch, no workspace. */
typedef struct GpFlowBnafDesc {
  int32_t abi_version;     /* GPFLOW_ABI_VERSION */
  int32_t dtype;           /* GPFLOW_F32 */
  int32_t dim;
  int32_t hidden_per_dim;  /* n_neurons / dim */
  int32_t n_layers;        /* masked linear layers, >= 2 */
  int32_t reserved0;
} GpFlowBnafDesc;
int gpflow_bnaf_param_count(const GpFlowBnafDesc* desc, int64_t* count);
int gpflow_bnaf_forward(const GpFlowBnafDesc* desc, int64_t B, const void* params, const void* x, void* y_out,
                        void* logdet_out, void* cuda_stream);

/* The frozen FFJORD map either side of the GP flow: `FFJORD.forward(x) -> (z, log_det)` --
 * `x_map, log_det = self.map(x)` of `Network.loss_gp_flow` (src/gp_flows/network.py:76-78, flow_models/ffjord/ffjord.py:50-53)
 * -- and `FFJORD.backward(z) -> x` (ffjord.py:55-56), the `map.backward` that builds the Gaussian-noise training set
 * (src/data_module.py:161-186).  The map is a chain of `n_blocks` CNF blocks (layers/container.py:14-26, layers/cnf.py:44-98)
 * whose vector field is an ODEnet of `n_hidden + 1` ConcatSquashLinear layers with tanh in between
 * (layers/odefunc.py:160-174, layers/diffeq_layers/basic.py:76-86); log_det integrates the exact trace of df/dy
 * (`divergence_bf`, odefunc.py:13-17).  Each block is integrated over [0, t_end[b]] (reverse: blocks in reverse order,
 * each over [t_end[b], 0]) with `n_steps` classical RK4 steps; the reference's adaptive torchdiffeq dopri5 is not
 * reproduced (see oracle/ffjord_oracle.py).
 * `params` = for each block, for each layer: `_layer.weight [out][in]`, `_layer.bias [out]`, `_hyper_bias.weight [out]`,
 * `_hyper_gate.weight [out]`, `_hyper_gate.bias [out]` (state_dict order).  Writes y_out [B, dim] and, if not NULL,
 * logdet_out [B] (log|det dy/dx| of the direction integrated).  One launch, no workspace.  Shapes whose per-thread
 * state does not fit in shared memory return GPFLOW_ERR_UNSUPPORTED. */
#define GPFLOW_FFJORD_MAX_HIDDEN 7
#define GPFLOW_FFJORD_MAX_BLOCKS 8
typedef struct GpFlowFfjordDesc {
  int32_t abi_version;     /* GPFLOW_ABI_VERSION */
  int32_t dtype;           /* GPFLOW_F32 / GPFLOW_F64 */
  int32_t dim;
  int32_t n_hidden;        /* hidden layers of the ODEnet, 0..GPFLOW_FFJORD_MAX_HIDDEN */
  int32_t hidden[GPFLOW_FFJORD_MAX_HIDDEN];
  int32_t n_blocks;        /* CNF blocks, 1..GPFLOW_FFJORD_MAX_BLOCKS */
  int32_t n_steps;         /* RK4 steps per block */
  int32_t reverse;         /* 0: FFJORD.forward, 1: FFJORD.backward */
  double t_end[GPFLOW_FFJORD_MAX_BLOCKS]; /* sqrt_end_time^2 of each block (cnf.py:49-52) */
} GpFlowFfjordDesc;
int gpflow_ffjord_param_count(const GpFlowFfjordDesc* desc, int64_t* count);
int gpflow_ffjord_integrate(const GpFlowFfjordDesc* desc, int64_t B, const void* params, const void* x, void* y_out,
                            void* logdet_out, void* cuda_stream);

const char* gpflow_error_string(int status);
int gpflow_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GPFLOW_B200_H */
