// Explicit instantiations of the fused stage kernel: variable-coefficient advection (scalar/varadv.py).
#include "dg1d_scalar_dispatch.cuh"

namespace dg1d {

int launch_stage_varadv(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.flux) {
    case DG1D_FLUX_UPWIND: return sc_by_degree<DG1D_PDE_VARADV, DG1D_FLUX_UPWIND>(h, a, st);
    }
    set_error("varadv: the numerical flux is the upwind form of varadv.py:11-13 (got %d)", h.cfg.flux);
    return DG1D_EINVAL;
}

}  // namespace dg1d
