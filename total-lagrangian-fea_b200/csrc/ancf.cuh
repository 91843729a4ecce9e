// ancf.cuh -- ANCF3243 beam / ANCF3443 shell specific set-up (shape-coefficient matrices,
// reference-gradient precompute, element mass).  The force/tangent kernels are shared (kernels.cuh).
#pragma once
#include "common.cuh"

namespace tlfea {
int ancf_setup(tlfea_elem *h, const double *L, const double *W, const double *H, const int *conn);
int ancf_precompute(tlfea_elem *h);
int ancf_elem_mass(tlfea_elem *h, double *d_Me);
}  // namespace tlfea
