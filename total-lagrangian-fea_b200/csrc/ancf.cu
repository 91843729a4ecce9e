#include "ancf.cuh"
namespace tlfea {
int ancf_setup(tlfea_elem *, const double *, const double *, const double *, const int *) {
  return fail(TLFEA_ERR_UNSUPPORTED, "ANCF set-up not built yet");
}
int ancf_precompute(tlfea_elem *) { return fail(TLFEA_ERR_UNSUPPORTED, "ANCF precompute not built yet"); }
int ancf_elem_mass(tlfea_elem *, double *) { return fail(TLFEA_ERR_UNSUPPORTED, "ANCF mass not built yet"); }
}  // namespace tlfea
