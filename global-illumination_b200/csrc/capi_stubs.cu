// csrc/capi_stubs.cu -- entry points of include/gi_b200.h whose kernels are not written yet.  They fail loudly
// (non-zero status + message); nothing here computes on the CPU.
#include "context.h"
extern "C" {
int gi_trace_photons(GiContext *, const GiParams *, GiPhotonStats *) {
  gi::set_error("gi_trace_photons: not implemented in this build"); return 1;
}
}
