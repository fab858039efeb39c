/*
 * dg1d.h -- C ABI of libdg1d.so: the explicit RK-DG time-stepping hot path of cpraveen/fem's
 * dg1d/python solvers (scalar advection / Burgers, 1-D Euler; Legendre modal basis; SSP-RK3;
 * TVB-minmod limiter) as hand-written sm_100a CUDA.
 *
 * Plain C: pointers, sizes and a stream handle -- no torch types.  All `double*` state arguments
 * are DEVICE pointers owned by the caller (the Python host layer passes torch tensors' data_ptr);
 * the library never frees or retains them.  Every call is asynchronous on `stream`
 * (a cudaStream_t passed as void*; NULL = legacy default stream) unless stated otherwise.
 * Every function returns 0 on success or a negative DG1D_E* code; dg1d_last_error() gives the text.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with DG1D_ECUDA.
 *
 * Citations are to files under the reference tree's dg1d/python/ directory.
 *
 * Device state layout ("SoA, ghosted"):  doubles [nvar][nd][nprob][pitch], cell index fastest.
 *   nvar = 1 (linear, burger) or 3 (euler: rho, mom, ene);  nd = degree+1 modal coefficients;
 *   nprob = number of independent problems in the batch (1 unless ensemble mode);
 *   pitch = dg1d_pitch(h) >= ncell + 2*DG1D_GHOST, a multiple of 16 doubles.
 *   Local cell j (0 <= j < ncell) sits at offset j + DG1D_GHOST of its row; ghost cells j = -2,-1 and
 *   j = ncell, ncell+1 are the two neighbour cells on each side when that side is periodic (kept
 *   up to date by the kernels themselves) or a halo (filled by dg1d_halo_unpack); unused for copy BC.
 * The reference layout ("AoS") is one C-contiguous (ncell, nd) array per variable
 *   (scalar/dgsolver.py:102-104, euler/euler.py:151-159); here [nvar][nprob][ncell][nd].
 */
#ifndef DG1D_H
#define DG1D_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DG1D_GHOST 4      /* ghost slots per side in a row (the inner two are used) */
#define DG1D_MAX_DEGREE 4 /* polynomial degree 0..4 */

enum { DG1D_PDE_LINEAR = 0, DG1D_PDE_BURGER = 1, DG1D_PDE_EULER = 2, DG1D_PDE_VARADV = 3 /* scalar/varadv.py */ };
/* numerical fluxes: scalar/linadv.py:6-11, scalar/burger.py:7-18, euler/gas.py:79-89;
 * DG1D_FLUX_ROE for Euler is an extension (no dg1d/python counterpart): PDE::roe_flux of dg2d/models/euler/pde.h:209-285 in
 * one space dimension, bit-identical to that function. */
enum { DG1D_FLUX_CENTRAL = 0, DG1D_FLUX_UPWIND = 1, DG1D_FLUX_ROE = 2, DG1D_FLUX_GODUNOV = 3, DG1D_FLUX_RUSANOV = 4,
       /* Euler only: the flux family of the C sibling, dg1d/c/euler/src/flux.c:29-42 (LFFlux), :171-236 (ECUSPFlux),
        * :241-302 (HLLCFlux), :307-406 (AUSMDVFlux), :105-134 (LFCFlux); numbering of its dg.h:7-11 plus 4.
        * Bit-identical to that file compiled by gcc. */
       DG1D_FLUX_LF = 5, DG1D_FLUX_ECUSP = 6, DG1D_FLUX_HLLC = 7, DG1D_FLUX_AUSMDV = 8, DG1D_FLUX_LFC = 9 };
enum { DG1D_BC_PERIODIC = 0, DG1D_BC_COPY = 1, DG1D_BC_HALO = 2,
       /* scalar PDEs, the C++ twin's BCType (dg1d/scalar_legendre/dg.h:44, :367-413): the boundary face takes
        * the physical flux of a prescribed state (inflow) or of the cell's own trace (outflow = DG1D_BC_COPY, whose
        * numflux(u,u) is that flux for every scalar flux here).  Cells next to a closed end are not limited. */
       DG1D_BC_INFLOW = 3, DG1D_BC_OUTFLOW = DG1D_BC_COPY };
enum { DG1D_LIMITER_NONE = 0, DG1D_LIMITER_TVB = 1 };

enum {
    DG1D_OK = 0,
    DG1D_EINVAL = -1, /* bad argument / unsupported combination */
    DG1D_ECUDA = -2,  /* CUDA runtime error (text in dg1d_last_error) */
    DG1D_ENOMEM = -3,
    DG1D_ESTATE = -4  /* call order (e.g. operators not set) */
};

typedef struct dg1d_handle dg1d_handle;

/* Problem/scheme description: the constructor arguments of scalar/dgsolver.py:11-65 (DG.__init__)
 * and the flags of euler/euler.py:16-47, reduced to what the kernels need. */
typedef struct dg1d_config {
    int32_t pde;      /* DG1D_PDE_* */
    int32_t flux;     /* DG1D_FLUX_* already resolved as dgsolver.py:43-59 does */
    int32_t degree;   /* k */
    int32_t limiter;  /* DG1D_LIMITER_*; scalar: `limit == 'yes'` (dgsolver.py:213); euler: k > 0 (euler.py:239) */
    int32_t char_lim; /* euler.py:256 */
    int32_t bc_left;  /* DG1D_BC_* of the local block's left end */
    int32_t bc_right;
    int32_t nprob;    /* batch of independent problems (ensemble mode), 1 .. 65535 */
    int64_t ncell;        /* cells per problem held locally (without ghosts) */
    int64_t cell_offset;  /* global index of local cell 0 (domain decomposition), else 0 */
    int64_t ncell_global; /* cells of the whole mesh (>= ncell) */
    double gamma;         /* euler/gas.py:3 (1.4) */
} dg1d_config;

/* Library/runtime probes (usable without a GPU). */
int dg1d_version(void);
const char *dg1d_last_error(void);
int dg1d_device_count(void); /* number of CUDA devices, 0 if none / no driver */

/* Handle = validated config + operator constants (host memory only).  The current CUDA device at
 * call time of the compute functions is used. */
int dg1d_create(const dg1d_config *cfg, dg1d_handle **out);
void dg1d_destroy(dg1d_handle *h);

/* Operator constants exactly as the reference's own Python expressions produce them:
 * Vf[q*nd+j] = phi_j(xg_q), Vg[i*nd+q] = phi_i'(xg_q)*wg_q (dgsolver.py:74-78, euler.py:56-59),
 * bm/bp = phi_j(-1)/phi_j(+1) (dgsolver.py:87-90), wg = Gauss weights, ark/brk (dgsolver.py:7-8),
 * sqrt3/isqrt3 (euler.py:11-12).  Host pointers; copied.  Checked, because the kernels rely on it: bm[0] = bp[0] = 1 and
 * row 0 of Vg is zero (orthonormal Legendre basis), ark[0] = 0 and brk[0] = 1 (the first stage is u - lam*res). */
int dg1d_set_operators(dg1d_handle *h, const double *Vf, const double *Vg, const double *bm, const double *bp,
                       const double *wg, const double ark[3], const double brk[3], double sqrt3, double isqrt3);

/* Mesh data for x-dependent fluxes (DG1D_PDE_VARADV, scalar/varadv.py:3-17; required for it, ignored otherwise):
 * xmin, dx as in dgsolver.py:37-39, xg = the k+1 Gauss nodes (dgsolver.py:71, host pointer), speed_const = the
 * constant max_speed of varadv.py:15-17 as the host's numpy evaluates it. */
int dg1d_set_mesh(dg1d_handle *h, double xmin, double dx, const double *xg, double speed_const);

/* Prescribed boundary states for DG1D_BC_INFLOW (scalar_legendre/dg.h:372-376, :397-401: boundary_condition at the stage
 * time, evaluated by the host as that code evaluates its Function).  `values`: DEVICE doubles, {left, right} per stage,
 * `stride` doubles apart (0: one pair for all stages); not copied, must stay valid while launches use it.  dg1d_stage,
 * dg1d_stage_part and dg1d_residual read pair 0; dg1d_run_steps reads pair 3*s + rk for stage rk of its s-th step, so
 * a table of the stage times t, t+dt, t+dt/2 of every step drives a whole run without host round trips. */
int dg1d_set_boundary(dg1d_handle *h, const double *values, int64_t stride);

int64_t dg1d_pitch(const dg1d_handle *h);         /* doubles per row */
int64_t dg1d_state_doubles(const dg1d_handle *h); /* nvar*nd*nprob*pitch */
int dg1d_nvar(const dg1d_handle *h);

/* Layout conversion, device to device.  aos: [nvar][nprob][ncell][nd].  dg1d_pack_state also
 * refreshes periodic ghosts. */
int dg1d_pack_state(const dg1d_handle *h, const double *aos, double *soa, void *stream);
int dg1d_unpack_state(const dg1d_handle *h, const double *soa, double *aos, void *stream);
/* The same for cells cell0 .. cell0+n-1 only; aos_chunk: [nvar][nprob][n][nd].  Lets a host state (the reference's arrays,
 * dgsolver.py:102-104) stream in and out through a small staging buffer, copy of one chunk overlapping the conversion of
 * the one before (Dg1dDevice.upload / download), instead of needing a second full-size array on the device. */
int dg1d_pack_cells(const dg1d_handle *h, const double *aos_chunk, double *soa, int64_t cell0, int64_t n, void *stream);
int dg1d_unpack_cells(const dg1d_handle *h, const double *soa, double *aos_chunk, int64_t cell0, int64_t n, void *stream);

/* Initial condition by L2 projection (dgsolver.py:107-122, euler.py:162-169): vals holds the IC
 * evaluated by the host at the Gauss points, [nvar][nprob][ncell][Nq]; u[i,j] = 0.5*vals[i,:].dot(Vf[:,j]*wg). */
int dg1d_project(const dg1d_handle *h, const double *vals, double *soa, void *stream);

/* CFL reduction (linadv.py:13, burger.py:21-22, gas.py:27-31): speed[p] = max over cells of the
 * cell-average wave speed.  speed: device [nprob]; overwritten. */
int dg1d_max_speed(const dg1d_handle *h, const double *soa, double *speed, void *stream);

/* Device-resident step control = dgsolver.py:263-268,275 / euler.py:181-186,274, one thread per problem:
 *   if done or !(t < Tf): done = 1; else dt = cfl*dx/speed; lam = dt/dx; if t+dt > Tf: dt = Tf-t, lam = dt/dx;
 *   dt_log[it] = dt (if it < dt_cap); t += dt; it += 1; speed = 0 (re-accumulated by stage rk=2).
 * All arrays are device [nprob] (dt_log: [nprob][dt_cap], may be NULL).  max_steps < 0: no step cap. */
int dg1d_step_control(const dg1d_handle *h, double cfl, double dx, double Tf, int64_t max_steps, double *t, double *dt,
                      double *lam, double *speed, int64_t *it, int32_t *done, double *dt_log, int64_t dt_cap,
                      void *stream);

/* One fused SSP-RK3 stage = residual (dgsolver.py:170-203 / euler.py:192-233) + update
 * (dgsolver.py:272 / euler.py:235-237) + limiter (dgsolver.py:206-228 / euler.py:239-273):
 *   u_out = limiter( ark[rk]*u_n + brk[rk]*(u_in - lam*res(u_in)) ).
 * u_in, u_n, u_out: three distinct state buffers (u_n is ignored for rk = 0, where u_n == u_in).
 * lam: device [nprob].  done: device [nprob] or NULL; problems with done != 0 are passed through
 * unchanged (rk = 0 copies u_in to u_out, rk = 1,2 leave u_out alone).  speed: device [nprob] or
 * NULL; when given, the max wave speed of u_out's cell averages is atomically max-ed into it. */
int dg1d_stage(const dg1d_handle *h, int rk, const double *u_in, const double *u_n, double *u_out, const double *lam,
               const int32_t *done, double Mdx2, double *speed, void *stream);

/* The same stage on a subset of the tiles, so that a domain-decomposed run can overlap the halo exchange with
 * arithmetic: DG1D_PART_EDGES = the first and last tile of every problem (they hold the cells the
 * neighbouring ranks wait for), DG1D_PART_INTERIOR = all other tiles.  EDGES followed by INTERIOR, in any order and on
 * any streams, equals one dg1d_stage call bit for bit. */
enum { DG1D_PART_EDGES = 1, DG1D_PART_INTERIOR = 2 };
/* part | DG1D_PART_DEPTH(d), d = 0, 1, 2: the edge part is d tiles wider on each side (and the interior part that much
 * narrower).  A tile next to an edge tile reads two of its cells; taking the edges of stage rk at depth rk, the interior
 * part of every stage of a step reads only what the interior part of the stage before wrote, so its three launches need
 * not wait for the edge tiles (dg1d_dist_run_steps does this).  The identity above holds at every depth. */
#define DG1D_PART_DEPTH(d) ((d) << 4)
int dg1d_stage_part(const dg1d_handle *h, int part, int rk, const double *u_in, const double *u_n, double *u_out,
                    const double *lam, const int32_t *done, double Mdx2, double *speed, void *stream);

/* Unfused pieces, for per-kernel parity tests against the reference's own functions. */
int dg1d_residual(const dg1d_handle *h, const double *u_in, double *res, void *stream);
int dg1d_rk_update(const dg1d_handle *h, int rk, const double *u_in, const double *u_n, const double *res,
                   const double *lam, double *u_out, void *stream);
int dg1d_limiter(const dg1d_handle *h, const double *u_in, double *u_out, double Mdx2, void *stream);

/* Time loop without host round trips: runs `nsteps` x (step_control + 3 stages) on `stream`.
 * bufs[0] holds the current state on entry and on return (the three pointers are rotated in place).
 * Requires bc PERIODIC/COPY (no halo).  Problems that reach Tf are passed through. */
int dg1d_run_steps(const dg1d_handle *h, int64_t nsteps, double *bufs[3], double cfl, double dx, double Tf,
                   int64_t max_steps, double Mdx2, double *t, double *dt, double *lam, double *speed, int64_t *it,
                   int32_t *done, double *dt_log, int64_t dt_cap, void *stream);

/* Domain decomposition halo: the two outermost cells per side, all modes and variables.
 * send_left/right, recv_left/right: device [nvar][nd][nprob][2] doubles each (dg1d_halo_doubles). NULL = skip side. */
int64_t dg1d_halo_doubles(const dg1d_handle *h);
int dg1d_halo_pack(const dg1d_handle *h, const double *soa, double *send_left, double *send_right, void *stream);
int dg1d_halo_unpack(const dg1d_handle *h, double *soa, const double *recv_left, const double *recv_right,
                     void *stream);

/* Domain-decomposed time loop driven from C (pattern of dg2d/system_legendre_mpi/dg.h:1026-1033, made
 * bit-identical to the serial order by the two-cell halo): nsteps x [step control; per stage: edge tiles, halo
 * exchange (ncclSend/ncclRecv on the handle's own stream) overlapped with the interior tiles; all-reduce(MAX) of the
 * wave speed].  One dg1d_dist per process and GPU; NCCL is taken from the libnccl.so.2 already loaded in the process.
 * left/right: neighbour ranks, -1 at a physical boundary (periodic meshes form a ring).  id: 128 bytes from
 * dg1d_dist_unique_id on one rank, distributed by the caller.  send/recv buffers: device, dg1d_halo_doubles(h) each.
 * `speed` must hold the globally reduced wave speed on entry (dg1d_max_speed + dg1d_dist_reduce_speed). */
typedef struct dg1d_dist dg1d_dist;
int dg1d_dist_available(void);
int dg1d_dist_unique_id(char id[128]);
int dg1d_dist_create(const char id[128], int world, int rank, int left, int right, dg1d_dist **out);
void dg1d_dist_destroy(dg1d_dist *d);
void dg1d_dist_invalidate_halo(dg1d_dist *d); /* after the state was replaced from outside */
int dg1d_dist_reduce_speed(dg1d_dist *d, double *speed, int nprob, void *stream);
/* Peer-memory transport (optional; single node, every pair of GPUs peer-accessible): each rank opens a mailbox in its
 * own HBM and exports it as a 64-byte CUDA IPC handle; after the caller has gathered all `world` handles (rank order)
 * dg1d_dist_p2p_connect maps them, and from then on the halo exchange and the CFL reduction of this dg1d_dist are
 * single kernels that store straight into the neighbours' memory over NVLink (no NCCL call in the step loop).  All ranks
 * must connect (or none), and synchronise (a barrier) between dg1d_dist_p2p_open and the first step so that no rank
 * writes into a mailbox that is not yet zeroed.  dg1d_dist_p2p_error synchronises the device and reports whether a
 * wait for a neighbour exceeded 20 s (the kernels give up instead of hanging).  Same results as NCCL, bit for bit. */
enum { DG1D_TRANSPORT_NCCL = 0, DG1D_TRANSPORT_PEER = 1 };
int dg1d_dist_p2p_open(dg1d_dist *d, const dg1d_handle *h, char ipc_handle[64]);
int dg1d_dist_p2p_connect(dg1d_dist *d, const char *ipc_handles /* [world][64] */);
int dg1d_dist_use_nccl(dg1d_dist *d); /* back to NCCL, e.g. when another rank could not connect (collective decision) */
int dg1d_dist_transport(const dg1d_dist *d);
int dg1d_dist_p2p_error(dg1d_dist *d, int64_t *timed_out);
int dg1d_dist_run_steps(const dg1d_handle *h, dg1d_dist *d, int64_t nsteps, double *bufs[3], double cfl, double dx, double Tf,
                        int64_t max_steps, double Mdx2, double *t, double *dt, double *lam, double *speed, int64_t *it,
                        int32_t *done, double *dt_log, int64_t dt_cap, double *send_left, double *send_right,
                        double *recv_left, double *recv_right, void *stream);

/* Host-buffer convenience (the reference-facing call: numpy arrays in, numpy arrays out).
 * u_aos: HOST [nvar][nprob][ncell][nd], overwritten with the state at Tf.  Allocates and frees its
 * own device memory; synchronous.  t_out/it_out: HOST [nprob] (may be NULL). */
int dg1d_solve_host(const dg1d_handle *h, double *u_aos, double cfl, double dx, double Tf, int64_t max_steps,
                    double Mdx2, double *t_out, int64_t *it_out);

/* L2 error norm on the device (scalar/dgsolver.py:232-252, scalar solver only): out[p] = sum over cells of
 * 0.5*dx*sum_q (V[q,:].u[i,:] - exact[p][i][q])^2 * w[q]  (the caller takes the square root, dgsolver.py:251).
 * V: HOST [ng][nd] = phi_j at the ng-point Gauss nodes, w: HOST [ng] weights (the reference uses ng = degree+3);
 * exact: DEVICE [nprob][ncell][ng], the exact solution at those points (evaluated by the host, like the IC);
 * scratch: DEVICE, dg1d_error_norm_scratch_doubles(h) doubles (per-block partial sums: the sum is done in two
 * passes in a fixed order, so it is reproducible); out: DEVICE [nprob]. */
int64_t dg1d_error_norm_scratch_doubles(const dg1d_handle *h);
int dg1d_error_norm(const dg1d_handle *h, const double *soa, const double *exact, const double *V, const double *w, int ng,
                    double dx, double *scratch, double *out, void *stream);

/* Uniform-point evaluation on the device: the numbers the reference computes for its plots (the display is left out).
 * Scalar (scalar/dgsolver.py:125-167): per cell f = Vu @ u1[i,:] at the nu points of xu = linspace(-1,1,nu),
 * nu = max(2, degree+1) (dgsolver.py:82-84), with the running umin/umax that become the axis limits (:132, :138-139).
 * Euler (euler/euler.py:77-148): density r = Vu.dot(rho[i,:]), velocity m/r, pressure(r, m, E) (gas.py:7-8) and the
 * running min/max of each.  Bit-identical to the reference for degree <= 2 (the small gemv's fused order; degree >= 3
 * agrees to 1 ulp), IEEE quotient, one rounding per operation in the pressure.
 * Vu: HOST [nu][nd], 1 <= nu <= 8.  out: DEVICE [nvar][nprob][ncell][nu] -- scalar: f; Euler: density, velocity, pressure.
 * minmax: DEVICE [nvar][nprob][2] = (min, max) per quantity and problem, started from (1e20, -1e20) as the reference
 * starts them, NaN-propagating as numpy's min/max are; may be NULL.  No scratch memory: the reduction runs on `minmax`
 * itself (it holds order-preserving integer keys between the three launches of this call). */
int dg1d_eval_points(const dg1d_handle *h, const double *soa, const double *Vu, int nu, double *out, double *minmax, void *stream);

/* dg1d_error_norm for callers that cannot allocate device memory.  It has no counterpart in the reference.
 * Who needs it: a monitor that evaluates the error during a run whose state buffers already fill the GPU (a
 * cudaMalloc could fail there and would stall the device), or a host that has handed all device memory to its own
 * pool and may not allocate behind it.  Same arguments and bit-identical result as dg1d_error_norm, minus
 * `scratch`: the per-block partial sums go to an array built into the library (64 KiB per GPU, held until exit).
 * Limit: dg1d_error_norm_scratch_doubles(h) <= dg1d_error_norm_noalloc_max_doubles() (= 8192, e.g. one problem of
 * any size, or 6 problems of >= 3e5 cells); larger requests fail with DG1D_EINVAL and a message, nothing is run.
 * The call allocates and frees nothing.  The first call on each device looks up the array's device address
 * (once, with that device current): make that first call during set-up, before starting any stream capture -- with
 * lazy module loading it can also be the moment the CUDA runtime loads this library's device code on that GPU, which is
 * the runtime's own (one-time) allocation.  The address is kept for the life of the process: a cudaDeviceReset
 * invalidates it, so resetting a device on which this entry point has been used is not supported.
 * All calls in the process share the one array per GPU: two calls whose kernels may overlap on the GPU (different
 * streams, no ordering between them) race on it and return garbage -- order them, or use dg1d_error_norm.
 * Nothing needs clearing between calls: every slot that is read was written earlier in the same call. */
int64_t dg1d_error_norm_noalloc_max_doubles(void);
int dg1d_error_norm_noalloc(const dg1d_handle *h, const double *soa, const double *exact, const double *V, const double *w,
                            int ng, double dx, double *out, void *stream);

/* Self-test hook (no reference counterpart): compares the kernels' shared-reciprocal division with the
 * compiler's IEEE a/b on n hashed operand pairs, bit for bit.  spread = half-width of the exponent range.
 * counts[0] = mismatches on the fast path (must be 0), counts[1] = pairs sent to the literal fallback,
 * counts[2] = exact-zero numerators generated.  Synchronous. */
int dg1d_selftest_division(int64_t n, uint64_t seed, int spread, uint64_t counts[3]);

/* Counter of kernels launched by this library in this process (for bench.py's gpu_launches). */
int64_t dg1d_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* DG1D_H */
