// Explicit instantiations of the fused stage kernel: 1-D Euler (euler/gas.py, euler/euler.py), Rusanov and Roe fluxes;
// the C sibling's flux family is instantiated by dg1d_euler_flux.cu, once per flux.
#include "dg1d_euler_dispatch.cuh"

namespace dg1d {

int launch_stage_euler_lf(const Handle &, const StageArgs &, cudaStream_t);
int launch_stage_euler_ecusp(const Handle &, const StageArgs &, cudaStream_t);
int launch_stage_euler_hllc(const Handle &, const StageArgs &, cudaStream_t);
int launch_stage_euler_ausmdv(const Handle &, const StageArgs &, cudaStream_t);
int launch_stage_euler_lfc(const Handle &, const StageArgs &, cudaStream_t);

int launch_stage_euler(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.flux) {
    case DG1D_FLUX_RUSANOV: return eu_by_degree<DG1D_FLUX_RUSANOV>(h, a, st);
    case DG1D_FLUX_ROE: return eu_by_degree<DG1D_FLUX_ROE>(h, a, st);
    case DG1D_FLUX_LF: return launch_stage_euler_lf(h, a, st);
    case DG1D_FLUX_ECUSP: return launch_stage_euler_ecusp(h, a, st);
    case DG1D_FLUX_HLLC: return launch_stage_euler_hllc(h, a, st);
    case DG1D_FLUX_AUSMDV: return launch_stage_euler_ausmdv(h, a, st);
    case DG1D_FLUX_LFC: return launch_stage_euler_lfc(h, a, st);
    }
    set_error("euler: numerical flux %d not available (rusanov, roe, lf, ecusp, hllc, ausmdv, lfc)", h.cfg.flux);
    return DG1D_EINVAL;
}

}  // namespace dg1d
