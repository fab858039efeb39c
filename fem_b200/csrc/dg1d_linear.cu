// Explicit instantiations of the fused stage kernel: scalar linear advection (scalar/linadv.py).
#include "dg1d_scalar_dispatch.cuh"

namespace dg1d {

int launch_stage_linear(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.flux) {
    case DG1D_FLUX_CENTRAL: return sc_by_degree<DG1D_PDE_LINEAR, DG1D_FLUX_CENTRAL>(h, a, st);
    case DG1D_FLUX_UPWIND: return sc_by_degree<DG1D_PDE_LINEAR, DG1D_FLUX_UPWIND>(h, a, st);
    }
    set_error("linear: numerical flux %d not available (central, upwind)", h.cfg.flux);
    return DG1D_EINVAL;
}

}  // namespace dg1d
