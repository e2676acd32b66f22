// csrc/capi_stubs.cu -- entry points of include/gi_b200.h whose kernels are not written yet.  They fail loudly
// (non-zero status + message); nothing here computes on the CPU.
#include "context.h"
extern "C" {
int gi_trace_photons(GiContext *, const GiParams *, GiPhotonStats *) {
  gi::set_error("gi_trace_photons: not implemented in this build"); return 1;
}
int gi_render(GiContext *, const GiParams *, int, int, int, uint8_t *, GiRenderStats *) {
  gi::set_error("gi_render: not implemented in this build"); return 1;
}
int gi_render_float(GiContext *, const GiParams *, int, int, int, double *, GiRenderStats *) {
  gi::set_error("gi_render_float: not implemented in this build"); return 1;
}
}
