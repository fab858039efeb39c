// Explicit instantiations of the fused stage kernel: inviscid Burgers (scalar/burger.py).
#include "dg1d_scalar_dispatch.cuh"

namespace dg1d {

int launch_stage_burger(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.flux) {
    case DG1D_FLUX_CENTRAL: return sc_by_degree<DG1D_PDE_BURGER, DG1D_FLUX_CENTRAL>(h, a, st);
    case DG1D_FLUX_ROE: return sc_by_degree<DG1D_PDE_BURGER, DG1D_FLUX_ROE>(h, a, st);
    case DG1D_FLUX_GODUNOV: return sc_by_degree<DG1D_PDE_BURGER, DG1D_FLUX_GODUNOV>(h, a, st);
    }
    set_error("burger: numerical flux %d not available (central, roe, godunov)", h.cfg.flux);
    return DG1D_EINVAL;
}

}  // namespace dg1d
