// translation unit: sf2 (2-D stream-function) kernels, float: train step / backward (TMA-stashed) and forward-only
#include "gpflow_sf2r_fwd.cuh"
#include "gpflow_sf2e_kernel.cuh"
namespace gpflow {
// GPFLOW_SF2_FORM=0 selects the first formulation (forward-mode tangents, 72-float stash: gpflow_sf2_kernel) for A/B
// measurements and as an independent cross-check; the default is the reverse-mode-gradient kernel gpflow_sf2r_kernel.
static int sf2_form() {
  static const int f = [] { const char* e = getenv("GPFLOW_SF2_FORM"); return (e != nullptr && e[0] == '0') ? 0 : 1; }();
  return f;
}
GPF_DECL_SMALL(sf2_15_15_f32) {
  const Ptrs<float>& p = *reinterpret_cast<const Ptrs<float>*>(ptrs);
  if (op == GPFLOW_OP_FORWARD)
    return sf2_form() == 0 ? launch_sf2_fwd<15, 15>(d, p, B, ws, wsb, st, sm_count) : launch_sf2r_fwd<15, 15>(d, p, B, ws, wsb, st, sm_count);
  if (sf2_form() == 0) return launch_sf2<15, 15>(d, op, p, B, ws, wsb, st, sm_count);
  return launch_sf2r<15, 15>(d, p, B, ws, wsb, st, sm_count);
}
// Euler penalization on the sf2r structure (gpflow_sf2e_kernel); GPFLOW_SF2_EULER=0 keeps it on the v1 dual kernel
bool sf2e_15_15_enabled() {
  static const int f = [] { const char* e = getenv("GPFLOW_SF2_EULER"); return (e != nullptr && e[0] == '0') ? 0 : 1; }();
  return f != 0 && sf2_form() != 0;
}
GPF_DECL_SMALL(sf2e_15_15_f32) {
  const Ptrs<float>& p = *reinterpret_cast<const Ptrs<float>*>(ptrs);
  if (op == GPFLOW_OP_FORWARD) return GPFLOW_ERR_UNSUPPORTED;
  return launch_sf2e<15, 15>(d, p, B, ws, wsb, st, sm_count);
}
size_t sf2e_15_15_workspace(const GpFlowDesc* d, int64_t B) { return sf2e_workspace<15, 15>(d, B); }
// kernels per call (launch accounting) and the name of the main kernel
int sf2_15_15_launches(int op) { (void)op; return (sf2_form() != 0 && GPF_SF2R_CONST) ? 3 : 2; }
const char* sf2_15_15_kernel_name(int op) {
  if (sf2_form() == 0) return "gpflow_sf2_kernel";
  return op == GPFLOW_OP_FORWARD ? "gpflow_sf2r_fwd_kernel" : "gpflow_sf2r_kernel";
}
size_t sf2_15_15_workspace(const GpFlowDesc* d, int64_t B, int op) {
  if (op == GPFLOW_OP_FORWARD) return sf2_form() == 0 ? sf2_fwd_workspace<15, 15>(B) : sf2r_fwd_workspace<15, 15>(B);
  return sf2_form() == 0 ? sf2_workspace<15, 15>(d, B) : sf2r_workspace<15, 15>(d, B);
}
}  // namespace gpflow
