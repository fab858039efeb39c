// Degree / limiter dispatch shared by the Euler translation units (one per numerical flux, so that they build in parallel).
#pragma once
#include "dg1d_kernels.cuh"

namespace dg1d {

template <int FLUX, int K>
static int eu_by_lim(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    // euler.py:239: the limiter runs whenever k > 0; char_lim picks characteristic variables (euler.py:256)
    if (h.cfg.limiter == DG1D_LIMITER_TVB && K > 0) {
        if (h.cfg.char_lim == 1) return launch_stage_t<DG1D_PDE_EULER, FLUX, K, (K > 0 ? 2 : 0)>(h, a, st);
        return launch_stage_t<DG1D_PDE_EULER, FLUX, K, (K > 0 ? 1 : 0)>(h, a, st);
    }
    return launch_stage_t<DG1D_PDE_EULER, FLUX, K, 0>(h, a, st);
}

template <int FLUX>
static int eu_by_degree(const Handle &h, const StageArgs &a, cudaStream_t st)
{
    switch (h.cfg.degree) {
    case 0: return eu_by_lim<FLUX, 0>(h, a, st);
    case 1: return eu_by_lim<FLUX, 1>(h, a, st);
    case 2: return eu_by_lim<FLUX, 2>(h, a, st);
    case 3: return eu_by_lim<FLUX, 3>(h, a, st);
    case 4: return eu_by_lim<FLUX, 4>(h, a, st);
    }
    set_error("degree %d not supported (0..%d)", h.cfg.degree, DG1D_MAX_DEGREE);
    return DG1D_EINVAL;
}

}  // namespace dg1d
