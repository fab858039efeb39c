// dg1d_dist.cu -- domain-decomposed time stepping driven from C: per stage a two-cell halo to each neighbour with
// ncclSend/ncclRecv on a communication stream, overlapped with the interior tiles; per step one ncclAllReduce(MAX)
// of the CFL wave speed.  The Python layer only creates the communicator and calls
// dg1d_dist_run_steps: at 0.1-0.5 ms per step the cost of issuing ~25 launches and three exchanges per step from
// Python was what limited the scalar PDEs on two GPUs.
//
// NCCL is resolved at run time from the libnccl.so.2 already in the process (PyTorch's), so libdg1d.so itself has
// no link-time dependency on it and still loads on a machine without NCCL.
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>
#include <new>

#include "dg1d_kernels.cuh"

namespace dg1d {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId *);
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char *(*GetErrorString)(ncclResult_t);
    bool ok;
};

static NcclApi *nccl_api()
{
    static NcclApi api = {};
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // PyTorch's copy, if it is in the process
        if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW);
        if (lib) {
#define SYM(field, name) api.field = (decltype(api.field))dlsym(lib, name)
            SYM(GetUniqueId, "ncclGetUniqueId");
            SYM(CommInitRank, "ncclCommInitRank");
            SYM(CommDestroy, "ncclCommDestroy");
            SYM(Send, "ncclSend");
            SYM(Recv, "ncclRecv");
            SYM(AllReduce, "ncclAllReduce");
            SYM(GroupStart, "ncclGroupStart");
            SYM(GroupEnd, "ncclGroupEnd");
            SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
            api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.Send && api.Recv && api.AllReduce &&
                     api.GroupStart && api.GroupEnd && api.GetErrorString;
        }
    }
    return api.ok ? &api : nullptr;
}

}  // namespace dg1d

using namespace dg1d;

struct dg1d_dist {
    ncclComm_t comm;
    int world, rank, left, right;  // neighbour ranks, -1 at a physical boundary
    cudaStream_t comm_stream;
    cudaEvent_t ev_edge, ev_int, ev_halo;
    const double *halo_valid;  // state buffer whose ghosts are up to date (nullptr: none)
};

#define NCCL_TRY(expr)                                                              \
    do {                                                                            \
        ncclResult_t r__ = (expr);                                                  \
        if (r__ != ncclSuccess) {                                                   \
            set_error("%s failed: %s", #expr, nccl_api()->GetErrorString(r__));     \
            return DG1D_ECUDA;                                                      \
        }                                                                           \
    } while (0)
#define CU_TRY(expr)                                                    \
    do {                                                                \
        cudaError_t e__ = (expr);                                       \
        if (e__ != cudaSuccess) {                                       \
            set_error("%s failed: %s", #expr, cudaGetErrorString(e__)); \
            return DG1D_ECUDA;                                          \
        }                                                               \
    } while (0)

// pack the two outermost cells per side of `soa`, exchange them with the neighbours, fill the ghosts of `soa`;
// everything on the communication stream
static int exchange(const dg1d_handle *h, dg1d_dist *d, double *soa, double *sl, double *sr, double *rl, double *rr)
{
    NcclApi *n = nccl_api();
    const size_t cnt = (size_t)dg1d_halo_doubles(h);
    const bool L = d->left >= 0, R = d->right >= 0;
    int rc = dg1d_halo_pack(h, soa, L ? sl : nullptr, R ? sr : nullptr, d->comm_stream);
    if (rc) return rc;
    NCCL_TRY(n->GroupStart());
    // my leftmost cells go to the left neighbour's right ghosts and vice versa.  On a two-rank ring both neighbours
    // are the same peer and NCCL pairs the n-th send with the peer's n-th receive: receive `right` first.
    if (L) NCCL_TRY(n->Send(sl, cnt, ncclDouble, d->left, d->comm, d->comm_stream));
    if (R) NCCL_TRY(n->Send(sr, cnt, ncclDouble, d->right, d->comm, d->comm_stream));
    if (R) NCCL_TRY(n->Recv(rr, cnt, ncclDouble, d->right, d->comm, d->comm_stream));
    if (L) NCCL_TRY(n->Recv(rl, cnt, ncclDouble, d->left, d->comm, d->comm_stream));
    NCCL_TRY(n->GroupEnd());
    return dg1d_halo_unpack(h, soa, L ? rl : nullptr, R ? rr : nullptr, d->comm_stream);
}

extern "C" {

int dg1d_dist_available(void) { return nccl_api() != nullptr; }

int dg1d_dist_unique_id(char id[128])
{
    NcclApi *n = nccl_api();
    if (!n) { set_error("libnccl.so.2 is not available in this process"); return DG1D_ESTATE; }
    if (!id) { set_error("null argument"); return DG1D_EINVAL; }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId u;
    NCCL_TRY(n->GetUniqueId(&u));
    memcpy(id, &u, 128);
    return DG1D_OK;
}

int dg1d_dist_create(const char id[128], int world, int rank, int left, int right, dg1d_dist **out)
{
    NcclApi *n = nccl_api();
    if (!n) { set_error("libnccl.so.2 is not available in this process"); return DG1D_ESTATE; }
    if (!id || !out || world < 2 || rank < 0 || rank >= world || left >= world || right >= world) {
        set_error("bad argument");
        return DG1D_EINVAL;
    }
    dg1d_dist *d = new (std::nothrow) dg1d_dist;
    if (!d) { set_error("out of host memory"); return DG1D_ENOMEM; }
    memset(d, 0, sizeof(*d));
    d->world = world; d->rank = rank; d->left = left; d->right = right;
    ncclUniqueId u;
    memcpy(&u, id, 128);
    NCCL_TRY(n->CommInitRank(&d->comm, world, u, rank));
    int prio_lo = 0, prio_hi = 0;
    CU_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CU_TRY(cudaStreamCreateWithPriority(&d->comm_stream, cudaStreamNonBlocking, prio_hi));  // edge tiles + exchange go first
    CU_TRY(cudaEventCreateWithFlags(&d->ev_edge, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&d->ev_int, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&d->ev_halo, cudaEventDisableTiming));
    *out = d;
    return DG1D_OK;
}

void dg1d_dist_destroy(dg1d_dist *d)
{
    if (!d) return;
    cudaStreamSynchronize(d->comm_stream);
    if (nccl_api() && d->comm) nccl_api()->CommDestroy(d->comm);
    cudaEventDestroy(d->ev_edge);
    cudaEventDestroy(d->ev_int);
    cudaEventDestroy(d->ev_halo);
    cudaStreamDestroy(d->comm_stream);
    delete d;
}

void dg1d_dist_invalidate_halo(dg1d_dist *d)
{
    if (d) d->halo_valid = nullptr;
}

int dg1d_dist_reduce_speed(dg1d_dist *d, double *speed, int nprob, void *stream)
{
    NcclApi *n = nccl_api();
    if (!n || !d || !speed) { set_error("bad argument"); return DG1D_EINVAL; }
    NCCL_TRY(n->AllReduce(speed, speed, (size_t)nprob, ncclDouble, ncclMax, d->comm, (cudaStream_t)stream));
    return DG1D_OK;
}

int dg1d_dist_run_steps(const dg1d_handle *h, dg1d_dist *d, int64_t nsteps, double *bufs[3], double cfl, double dx, double Tf,
                        int64_t max_steps, double Mdx2, double *t, double *dt, double *lam, double *speed, int64_t *it,
                        int32_t *done, double *dt_log, int64_t dt_cap, double *send_left, double *send_right,
                        double *recv_left, double *recv_right, void *stream)
{
    if (!h || !d || !bufs || !bufs[0] || !bufs[1] || !bufs[2] || !send_left || !send_right || !recv_left || !recv_right) {
        set_error("null argument");
        return DG1D_EINVAL;
    }
    NcclApi *n = nccl_api();
    cudaStream_t main = (cudaStream_t)stream, edge = d->comm_stream;
    int rc;
    // Two streams per stage.  `main`: the interior tiles.  `edge` (high priority): the first and last tile, then the
    // exchange of their results.  Both kernels of a stage run side by side; the exchange hides behind the interior.
    //   edges(s)    needs the ghosts of its input (exchange(s-1): same stream) and the interior tiles next to it
    //               (interior(s-1): ev_int);
    //   interior(s) needs interior(s-1) (same stream) and the edge tiles of stage s-1 (ev_edge).
    for (int64_t s = 0; s < nsteps; ++s) {
        if ((rc = dg1d_step_control(h, cfl, dx, Tf, max_steps, t, dt, lam, speed, it, done, dt_log, dt_cap, stream))) return rc;
        CU_TRY(cudaEventRecord(d->ev_int, main));  // lam/done of this step; everything queued so far on main
        double *X = bufs[0], *Y = bufs[1], *Z = bufs[2];
        const double *uin[3] = {X, Y, Z}, *un[3] = {nullptr, X, X};
        double *uout[3] = {Y, Z, Y};
        for (int rk = 0; rk < 3; ++rk) {
            double *sp = (rk == 2) ? speed : nullptr;
            CU_TRY(cudaStreamWaitEvent(edge, d->ev_int, 0));
            if (rk == 0 && d->halo_valid != X) {  // first step, or the state was replaced: fill the ghosts of u^n
                if ((rc = exchange(h, d, X, send_left, send_right, recv_left, recv_right))) return rc;
            }
            if ((rc = dg1d_stage_part(h, DG1D_PART_EDGES, rk, uin[rk], un[rk], uout[rk], lam, done, Mdx2, sp, edge))) return rc;
            CU_TRY(cudaEventRecord(d->ev_edge, edge));
            if ((rc = exchange(h, d, uout[rk], send_left, send_right, recv_left, recv_right))) return rc;
            if ((rc = dg1d_stage_part(h, DG1D_PART_INTERIOR, rk, uin[rk], un[rk], uout[rk], lam, done, Mdx2, sp, stream))) return rc;
            CU_TRY(cudaEventRecord(d->ev_int, main));
            CU_TRY(cudaStreamWaitEvent(main, d->ev_edge, 0));  // the next interior launch (or the reduction) needs the edge tiles
        }
        bufs[0] = Y; bufs[1] = X; bufs[2] = Z;
        d->halo_valid = Y;
        NCCL_TRY(n->AllReduce(speed, speed, (size_t)h->h.cfg.nprob, ncclDouble, ncclMax, d->comm, main));  // F9; max is exact
    }
    CU_TRY(cudaEventRecord(d->ev_halo, edge));
    CU_TRY(cudaStreamWaitEvent(main, d->ev_halo, 0));  // the caller's stream also sees the last ghosts
    return DG1D_OK;
}

}  // extern "C"
