// dg1d_norm.cu -- L2 error norm of the DG solution on the device (scalar/dgsolver.py:232-252).
//
// Reference: with ng = k+3 Gauss points per cell, error^2 = sum_i 0.5*dx*sum_q (Vf[q,:].u1[i,:] - ue[i,q])^2 * wg[q],
// accumulated cell after cell.  Here: one thread per cell forms the cell's term, a block adds its cells in a fixed
// tree order and writes ONE partial sum to scratch memory, and a second one-block kernel adds the partials in a
// fixed order.  No floating-point atomics: the result is reproducible run to run (it differs from the
// reference's left-to-right sum by summation order only, ~1e-16 relative).
#include <mutex>

#include "dg1d_kernels.cuh"

namespace dg1d {

constexpr int NORM_BLOCK = 256;
constexpr int NORM_MAX_BLOCKS = 1184;  // 148 SMs x 8; larger meshes use a grid-stride loop

struct NormOps {
    int ng, nd;
    double V[7 * 5];  // V[q*nd+j] = phi_j(xq), (k+3)-point rule
    double w[7];
};

__device__ __forceinline__ double block_sum(double v, double *sm)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) sm[warp] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += sm[w];
    return s;  // valid in thread 0
}

// partial[prob*gridDim.x + block] = sum over the block's cells of 0.5*dx*sum_q (un - ue)^2 wg
__global__ void norm_partial_kernel(NormOps o, i64 nc, i64 pitch, int nprob, double dx, const double *__restrict__ soa,
                                    const double *__restrict__ exact, double *__restrict__ partial)
{
    __shared__ double sm[NORM_BLOCK / 32];
    const int prob = blockIdx.y;
    const i64 rs = pitch * nprob;
    const double *row = soa + (i64)prob * pitch + GS;
    double acc = 0.0;
    for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < nc; j += (i64)gridDim.x * blockDim.x) {
        double u[5];
        for (int m = 0; m < o.nd; ++m) u[m] = row[m * rs + j];
        const double *ue = exact + ((i64)prob * nc + j) * o.ng;
        double e = 0.0;
        for (int q = 0; q < o.ng; ++q) {
            double un = o.V[q * o.nd] * u[0];
            for (int m = 1; m < o.nd; ++m) un = fma(o.V[q * o.nd + m], u[m], un);
            const double d = un - ue[q];
            e = e + (d * d) * o.w[q];
        }
        acc = acc + 0.5 * dx * e;
    }
    const double s = block_sum(acc, sm);
    if (threadIdx.x == 0) partial[(i64)prob * gridDim.x + blockIdx.x] = s;
}

// out[prob] = sum of the problem's partials, fixed order
__global__ void norm_final_kernel(int nblocks, const double *__restrict__ partial, double *__restrict__ out)
{
    __shared__ double sm[NORM_BLOCK / 32];
    const int prob = blockIdx.x;
    double acc = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) acc += partial[(i64)prob * nblocks + b];
    const double s = block_sum(acc, sm);
    if (threadIdx.x == 0) out[prob] = s;
}

static int norm_blocks(const Handle &h)
{
    i64 b = (h.cfg.ncell + NORM_BLOCK - 1) / NORM_BLOCK;
    return (int)(b < NORM_MAX_BLOCKS ? b : NORM_MAX_BLOCKS);
}

static int norm_launch(const Handle &h, const double *soa, const double *exact, const double *V, const double *w,
                       int ng, double dx, double *scratch, double *out, cudaStream_t st)
{
    NormOps o;
    o.ng = ng;
    o.nd = h.nd;
    for (int i = 0; i < ng * h.nd; ++i) o.V[i] = V[i];
    for (int i = 0; i < ng; ++i) o.w[i] = w[i];
    const int nb = norm_blocks(h);
    norm_partial_kernel<<<dim3((unsigned)nb, (unsigned)h.cfg.nprob), NORM_BLOCK, 0, st>>>(o, h.cfg.ncell, h.pitch, h.cfg.nprob, dx,
                                                                                      soa, exact, scratch);
    count_launch();
    norm_final_kernel<<<(unsigned)h.cfg.nprob, NORM_BLOCK, 0, st>>>(nb, scratch, out);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("error-norm kernels: %s", cudaGetErrorString(e));
        return DG1D_ECUDA;
    }
    return DG1D_OK;
}

// ---- built-in scratch for callers that must not allocate --------------------------------------------------------
// Some users of this library cannot hand dg1d_error_norm a scratch buffer: an in-situ monitor that evaluates the
// error every few steps of a run whose state already fills the card (a cudaMalloc there can fail, and it stalls the
// whole device while it runs), or a host program that has given every byte to its own memory pool and is not
// allowed to go behind it.  For them the per-block partial sums can live in this fixed array instead.
// NORM_STATIC_MAX doubles = 64 KiB; every GPU that loads libdg1d.so carries it until the process exits.
constexpr int NORM_STATIC_MAX = 8192;
static __device__ __align__(256) double g_norm_scratch[NORM_STATIC_MAX];

// In host code `g_norm_scratch` is only a handle for the symbol, not a device address; the address a kernel can use
// has to be asked for, once per device, while that device is current.
constexpr int NORM_MAX_DEVICES = 64;
static double *g_norm_scratch_addr[NORM_MAX_DEVICES] = {nullptr};
static std::mutex g_norm_mutex;

static int norm_static_scratch(double **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= NORM_MAX_DEVICES) {
        set_error("cannot identify the current device: %s", cudaGetErrorString(e));
        return DG1D_ECUDA;
    }
    std::lock_guard<std::mutex> lock(g_norm_mutex);
    if (g_norm_scratch_addr[dev] == nullptr) {
        double *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_norm_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            set_error("cudaGetSymbolAddress(built-in scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
            return DG1D_ECUDA;
        }
        g_norm_scratch_addr[dev] = ptr;
    }
    *out = g_norm_scratch_addr[dev];
    return DG1D_OK;
}

}  // namespace dg1d

using namespace dg1d;

static int norm_check(const dg1d_handle *h, const double *soa, const double *exact, const double *V, const double *w, int ng,
                      const double *out)
{
    if (!h || !soa || !exact || !V || !w || !out) {
        set_error("null argument");
        return DG1D_EINVAL;
    }
    if (h->h.nvar != 1) {
        set_error("the error norm is defined for the scalar solver only (dgsolver.py:232-252)");
        return DG1D_EINVAL;
    }
    if (ng < 1 || ng > 7) {
        set_error("ng must be 1..7 (the reference uses degree+3)");
        return DG1D_EINVAL;
    }
    return DG1D_OK;
}

extern "C" {

int64_t dg1d_error_norm_scratch_doubles(const dg1d_handle *h) { return h ? (int64_t)norm_blocks(h->h) * h->h.cfg.nprob : 0; }

int dg1d_error_norm(const dg1d_handle *h, const double *soa, const double *exact, const double *V, const double *w, int ng,
                    double dx, double *scratch, double *out, void *stream)
{
    int rc = norm_check(h, soa, exact, V, w, ng, out);
    if (rc) return rc;
    if (!scratch) {
        set_error("scratch must hold dg1d_error_norm_scratch_doubles(h) doubles");
        return DG1D_EINVAL;
    }
    return norm_launch(h->h, soa, exact, V, w, ng, dx, scratch, out, (cudaStream_t)stream);
}

int64_t dg1d_error_norm_noalloc_max_doubles(void) { return NORM_STATIC_MAX; }

// Same result as dg1d_error_norm, bit for bit (same two kernels, same summation order), with the partial sums in
// the library's built-in device array.  Nothing is allocated or freed here, by CUDA or by anyone else.
// Nothing in the array has to be zero beforehand: the first kernel overwrites every slot the second one reads,
// so what an earlier call left behind is never seen.
int dg1d_error_norm_noalloc(const dg1d_handle *h, const double *soa, const double *exact, const double *V, const double *w,
                            int ng, double dx, double *out, void *stream)
{
    int rc = norm_check(h, soa, exact, V, w, ng, out);
    if (rc) return rc;
    const int64_t need = dg1d_error_norm_scratch_doubles(h);
    if (need > NORM_STATIC_MAX) {
        set_error("dg1d_error_norm_noalloc: this problem needs %lld partial sums but the built-in scratch holds %d "
                  "(blocks per problem x problems); use dg1d_error_norm with your own scratch buffer",
                  (long long)need, NORM_STATIC_MAX);
        return DG1D_EINVAL;
    }
    double *scratch = nullptr;
    if ((rc = norm_static_scratch(&scratch))) return rc;
    return norm_launch(h->h, soa, exact, V, w, ng, dx, scratch, out, (cudaStream_t)stream);
}

}  // extern "C"
