// One flux of the C sibling's family (dg1d/c/euler/src/flux.c) per object file: built five times by the Makefile with
// -DDG1D_ONE_FLUX=<DG1D_FLUX_*> -DDG1D_ONE_NAME=<lf|ecusp|hllc|ausmdv|lfc>.
#include "dg1d_euler_dispatch.cuh"

#if !defined(DG1D_ONE_FLUX) || !defined(DG1D_ONE_NAME)
#error "build through the Makefile: -DDG1D_ONE_FLUX=... -DDG1D_ONE_NAME=..."
#endif
#define DG1D_CAT2(a, b) a##b
#define DG1D_CAT(a, b) DG1D_CAT2(a, b)

namespace dg1d {

int DG1D_CAT(launch_stage_euler_, DG1D_ONE_NAME)(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    return eu_by_degree<DG1D_ONE_FLUX>(h, a, st);
}

}  // namespace dg1d
