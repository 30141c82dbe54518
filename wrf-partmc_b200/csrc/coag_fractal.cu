// coag_fractal.cu -- the second build of coag.cu: mc_coag for fractal particles (pmcgpu_set_fractal) and for the
// brown_free / brown_cont kernels (coag_kernel_brown_free.F90, coag_kernel_brown_cont.F90).  See the note at the
// top of coag.cu: the spherical build stays free of this code.
#define PMC_COAG_FRACTAL 1
#include "coag.cu"
