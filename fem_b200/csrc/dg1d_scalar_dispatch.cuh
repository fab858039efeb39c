// Degree / limiter dispatch shared by the scalar translation units (one per PDE, so that they build in parallel).
#pragma once
#include "dg1d_kernels.cuh"

namespace dg1d {

template <int PDE, int FLUX, int K>
static int sc_by_lim(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    // dgsolver.py:213: the limiter runs when `limit == 'yes'` (degree 0 with the limiter is refused by dg1d_create)
    if (h.cfg.limiter == DG1D_LIMITER_TVB && K > 0) return launch_stage_t<PDE, FLUX, K, (K > 0 ? 1 : 0)>(h, a, st);
    return launch_stage_t<PDE, FLUX, K, 0>(h, a, st);
}

template <int PDE, int FLUX>
static int sc_by_degree(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.degree) {
    case 0: return sc_by_lim<PDE, FLUX, 0>(h, a, st);
    case 1: return sc_by_lim<PDE, FLUX, 1>(h, a, st);
    case 2: return sc_by_lim<PDE, FLUX, 2>(h, a, st);
    case 3: return sc_by_lim<PDE, FLUX, 3>(h, a, st);
    case 4: return sc_by_lim<PDE, FLUX, 4>(h, a, st);
    }
    set_error("degree %d not supported (0..%d)", h.cfg.degree, DG1D_MAX_DEGREE);
    return DG1D_EINVAL;
}

}  // namespace dg1d
