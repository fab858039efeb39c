// dg1d_eval.cu -- uniform-point evaluation of the solution on the device (second half).
//
// The reference evaluates every cell at nu = max(2, k+1) uniform points only to hand the values to matplotlib, together with
// their running minimum and maximum, which become the axis limits (scalar/dgsolver.py:125-167: f = Vu @ u1[i,:];
// euler/euler.py:77-148: r = Vu.dot(rho[i,:]), v = m/r, pre = pressure(r, m, E)).  The display is left out; the numbers
// are what a user of a 1e7-cell run needs in order to look at the solution without unpacking and copying the state.
//
// One thread per cell: coefficients from the SoA rows (coalesced), the nu point values of one quantity staged in shared
// memory and written as one contiguous run per block ([cell][nu] is the reference's order), min/max by warp shuffles and
// one atomic per block on an order-preserving 64-bit key.  Arithmetic as numpy's: the rows of Vu @ u in the fused order of
// the small gemv (mvrow, dg1d_kernels.cuh), m/r the IEEE quotient, the pressure one rounding per operation.
#include "dg1d_kernels.cuh"

namespace dg1d {

constexpr int EVAL_BLOCK = 128;
constexpr int EVAL_MAXNU = 8;

struct EvalParams {
    const double *soa;
    double *out;      // [nq][nprob][nc][nu]
    u64 *keys;        // [nq][nprob][2] order-preserving keys of (min, max), or null
    i64 nc, pitch;
    int nvar, nd, nu, nprob;
    double gamma;
    double Vu[EVAL_MAXNU * 5];  // [nu][nd]
};

// doubles -> unsigned keys with the same order (-inf < ... < -0 < +0 < ... < +inf)
__device__ __forceinline__ u64 ordered_key(double x)
{
    const u64 b = (u64)__double_as_longlong(x);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double key_value(u64 k)
{
    const u64 b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

__global__ void eval_init_kernel(u64 *keys, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = ordered_key((i & 1) ? -1.0e20 : 1.0e20);  // dgsolver.py:132: umin, umax = 1.0e20, -1.0e20
}

// keys -> doubles in place.  A NaN value was recorded as key 0 (minimum) / ~0 (maximum), which decode to NaN bit patterns:
// numpy's min and max propagate NaN, and once the running value is NaN it stays NaN.
__global__ void eval_final_kernel(u64 *keys, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) ((double *)keys)[i] = key_value(keys[i]);
}

template <int ND>
__device__ __forceinline__ double eval_row(const double *a, const double (&x)[ND])
{
    if (ND == 1) return a[0] * x[0];
    double t = fma(a[0], x[0], a[ND > 1 ? 1 : 0] * x[ND > 1 ? 1 : 0]);
#pragma unroll
    for (int i = 2; i < ND; ++i) t = fma(a[i], x[i], t);
    return t;
}

template <int NVAR, int ND>
__global__ void __launch_bounds__(EVAL_BLOCK) eval_points_kernel(const __grid_constant__ EvalParams p)
{
    __shared__ double stage[EVAL_BLOCK * EVAL_MAXNU];
    __shared__ u64 smin[EVAL_BLOCK / 32], smax[EVAL_BLOCK / 32];
    const int tid = threadIdx.x, prob = blockIdx.y;
    const i64 j0 = (i64)blockIdx.x * EVAL_BLOCK, j = j0 + tid;
    const bool live = j < p.nc;
    const i64 rs = p.pitch * p.nprob;
    constexpr int nd = ND;
    const int nu = p.nu;
    double u[NVAR][ND];
#pragma unroll
    for (int v = 0; v < NVAR; ++v)
#pragma unroll
        for (int m = 0; m < ND; ++m) u[v][m] = live ? p.soa[((i64)v * nd + m) * rs + (i64)prob * p.pitch + GS + j] : 0.0;
    const int ncell_blk = (int)((p.nc - j0 < EVAL_BLOCK) ? (p.nc - j0) : EVAL_BLOCK);
#pragma unroll
    for (int q = 0; q < NVAR; ++q) {
        u64 kmin = ~0ull, kmax = 0ull;
        for (int s = 0; s < nu; ++s) {
            double val;
            if constexpr (NVAR == 1) {
                val = eval_row<ND>(&p.Vu[s * nd], u[0]);                       // dgsolver.py:135
            } else {
                const double r = eval_row<ND>(&p.Vu[s * nd], u[0]);            // euler.py:86
                if (q == 0) {
                    val = r;
                } else {
                    const double m = eval_row<ND>(&p.Vu[s * nd], u[1]);        // euler.py:91
                    if (q == 1) val = m / r;                                   // euler.py:92
                    else val = eu_pressure(p.gamma, r, m, eval_row<ND>(&p.Vu[s * nd], u[2]));  // euler.py:97-98, gas.py:7-8
                }
            }
            stage[tid * nu + s] = val;
            if (live) {
                const bool nan = val != val;
                const u64 k = ordered_key(val);
                const u64 lo = nan ? 0ull : k, hi = nan ? ~0ull : k;
                kmin = lo < kmin ? lo : kmin;
                kmax = hi > kmax ? hi : kmax;
            }
        }
        __syncthreads();
        double *dst = p.out + (((i64)q * p.nprob + prob) * p.nc + j0) * nu;
        for (int i = tid; i < ncell_blk * nu; i += EVAL_BLOCK) dst[i] = stage[i];
        if (p.keys != nullptr) {
            for (int o = 16; o > 0; o >>= 1) {
                const u64 a = __shfl_xor_sync(0xffffffffu, kmin, o), b = __shfl_xor_sync(0xffffffffu, kmax, o);
                kmin = a < kmin ? a : kmin;
                kmax = b > kmax ? b : kmax;
            }
            if ((tid & 31) == 0) { smin[tid >> 5] = kmin; smax[tid >> 5] = kmax; }
        }
        __syncthreads();  // also: `stage` is free for the next quantity
        if (p.keys != nullptr && tid == 0) {
            for (int w = 1; w < EVAL_BLOCK / 32; ++w) {
                kmin = smin[w] < kmin ? smin[w] : kmin;
                kmax = smax[w] > kmax ? smax[w] : kmax;
            }
            atomicMin(p.keys + ((i64)q * p.nprob + prob) * 2, kmin);
            atomicMax(p.keys + ((i64)q * p.nprob + prob) * 2 + 1, kmax);
        }
    }
}

}  // namespace dg1d

using namespace dg1d;

extern "C" int dg1d_eval_points(const dg1d_handle *h, const double *soa, const double *Vu, int nu, double *out, double *minmax,
                                void *stream)
{
    if (!h || !soa || !Vu || !out) { set_error("null argument"); return DG1D_EINVAL; }
    if (nu < 1 || nu > EVAL_MAXNU) { set_error("nu must be 1..%d (the reference uses max(2, degree+1))", EVAL_MAXNU); return DG1D_EINVAL; }
    const Handle &H = h->h;
    EvalParams p;
    p.soa = soa; p.out = out; p.keys = (u64 *)minmax;
    p.nc = H.cfg.ncell; p.pitch = H.pitch; p.nvar = H.nvar; p.nd = H.nd; p.nu = nu; p.nprob = H.cfg.nprob;
    p.gamma = H.cfg.gamma;
    for (int i = 0; i < nu * H.nd; ++i) p.Vu[i] = Vu[i];
    cudaStream_t st = (cudaStream_t)stream;
    const int nkeys = H.nvar * H.cfg.nprob * 2;
    cudaError_t e = cudaSuccess;
    if (minmax) {
        eval_init_kernel<<<(nkeys + 127) / 128, 128, 0, st>>>(p.keys, nkeys);
        count_launch();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) {
        dim3 grid((unsigned)((H.cfg.ncell + EVAL_BLOCK - 1) / EVAL_BLOCK), (unsigned)H.cfg.nprob);
#define EVAL_CASE(NV, NDD)                                                     \
    case NDD:                                                                  \
        eval_points_kernel<NV, NDD><<<grid, EVAL_BLOCK, 0, st>>>(p);           \
        break;
        if (H.nvar == 1) {
            switch (H.nd) { EVAL_CASE(1, 1) EVAL_CASE(1, 2) EVAL_CASE(1, 3) EVAL_CASE(1, 4) EVAL_CASE(1, 5) }
        } else {
            switch (H.nd) { EVAL_CASE(3, 1) EVAL_CASE(3, 2) EVAL_CASE(3, 3) EVAL_CASE(3, 4) EVAL_CASE(3, 5) }
        }
#undef EVAL_CASE
        count_launch();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess && minmax) {
        eval_final_kernel<<<(nkeys + 127) / 128, 128, 0, st>>>(p.keys, nkeys);
        count_launch();
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        set_error("eval_points: %s", cudaGetErrorString(e));
        return DG1D_ECUDA;
    }
    return DG1D_OK;
}
