// Explicit instantiations of the fused stage kernel: variable-coefficient advection (scalar/varadv.py).
#include "dg1d_kernels.cuh"

namespace dg1d {

template <int FLUX, int K>
static int by_lim(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    if (h.cfg.limiter == DG1D_LIMITER_TVB && K > 0) return launch_stage_t<DG1D_PDE_VARADV, FLUX, K, (K > 0 ? 1 : 0)>(h, a, st);
    return launch_stage_t<DG1D_PDE_VARADV, FLUX, K, 0>(h, a, st);
}

template <int FLUX>
static int by_degree(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.degree) {
    case 0: return by_lim<FLUX, 0>(h, a, st);
    case 1: return by_lim<FLUX, 1>(h, a, st);
    case 2: return by_lim<FLUX, 2>(h, a, st);
    case 3: return by_lim<FLUX, 3>(h, a, st);
    case 4: return by_lim<FLUX, 4>(h, a, st);
    }
    set_error("degree %d not supported (0..%d)", h.cfg.degree, DG1D_MAX_DEGREE);
    return DG1D_EINVAL;
}

int launch_stage_varadv(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.flux) {
    case DG1D_FLUX_UPWIND: return by_degree<DG1D_FLUX_UPWIND>(h, a, st);
    }
    set_error("varadv: the numerical flux is the upwind form of varadv.py:11-13 (got %d)", h.cfg.flux);
    return DG1D_EINVAL;
}

}  // namespace dg1d
