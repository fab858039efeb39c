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

#include <cstddef>
#include <cstdlib>
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

// ---- peer-memory transport ---------------------------------------------------------------------------------------
// Every rank owns one "mailbox" in its HBM, mapped into the other ranks with CUDA IPC.  A halo exchange is ONE small
// kernel per rank: it stores the rank's two outermost cells per side straight into the neighbours' mailboxes over NVLink,
// publishes an epoch number behind a system-scope fence, waits for the neighbours' epoch in its own mailbox and fills
// its ghost slots -- no pack buffer, no NCCL kernel, no unpack launch.  The CFL reduction is the same pattern with
// every rank writing its wave speed into every mailbox.  Slots are double-buffered by epoch parity: a rank can only
// be one exchange ahead of a neighbour (it needs the neighbour's data of exchange e to finish e itself), so epoch e+2
// never overwrites data of epoch e that is still unread.  A wait that lasts longer than WAIT_LIMIT_NS gives up and
// raises the mailbox's error word instead of hanging the GPU.
constexpr int P2P_MAX_WORLD = 16;
constexpr unsigned long long WAIT_LIMIT_NS = 20ull * 1000ull * 1000ull * 1000ull;

struct MailboxHeader {  // at offset 0 of the mailbox; the rest follows at 256-byte aligned offsets
    unsigned long long halo_epoch[2];               // [0]: data from the left neighbour, [1]: from the right
    unsigned long long speed_epoch[P2P_MAX_WORLD];  // per source rank
    unsigned long long error;                       // non-zero: a wait timed out
    unsigned long long cell_epoch[4];               // [side][cell]: messages sent by the stage kernel itself (HaloLink)
};
static_assert(sizeof(MailboxHeader) <= 256, "the mailbox header must fit in front of the first slot");

struct MailboxGeom {
    long long halo_off, speed_off, bytes;  // byte offsets of halo[2 sides][2 parities][hd] and speed[2][world][nprob]
    int hd, world, nprob;
};

static MailboxGeom mailbox_geom(int hd, int world, int nprob)
{
    MailboxGeom g;
    g.hd = hd; g.world = world; g.nprob = nprob;
    g.halo_off = 256;
    g.speed_off = g.halo_off + (((long long)4 * hd * 8 + 255) / 256) * 256;
    g.bytes = g.speed_off + (((long long)2 * world * nprob * 8 + 255) / 256) * 256;
    return g;
}

__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// true when *flag reached `epoch`; false (and the error word raised) after WAIT_LIMIT_NS
__device__ __forceinline__ bool wait_epoch(const volatile unsigned long long *flag, unsigned long long epoch, MailboxHeader *mine)
{
    const volatile unsigned long long *err = &mine->error;
    const unsigned long long t0 = globaltimer_ns();
    while (*flag < epoch) {
        if (*err != 0ull) return false;  // an earlier wait already gave up: do not wait 20 s again at every exchange
        __nanosleep(64);
        if (globaltimer_ns() - t0 > WAIT_LIMIT_NS) {
            atomicExch(&mine->error, 1ull);
            return false;
        }
    }
    return true;
}

struct HaloGeo {
    long long nc, pitch;
    int rows;  // nvar*nd*nprob
};

// blockIdx.x = side (0: towards/from the left neighbour, 1: right).  halo slot layout as dg1d_halo_pack: [row][2].
__global__ void p2p_halo_exchange_kernel(HaloGeo g, MailboxGeom mg, double *soa, char *mine, char *peer_left, char *peer_right,
                                         unsigned long long epoch)
{
    const int side = blockIdx.x;
    char *peer = side == 0 ? peer_left : peer_right;
    if (peer == nullptr) return;  // physical boundary on this side
    const int n = g.rows * 2;
    const int par = (int)(epoch & 1ull);
    // my leftmost cells are the left neighbour's RIGHT ghosts (its side-1 slot) and vice versa
    double *dst = (double *)(peer + mg.halo_off) + ((long long)(1 - side) * 2 + par) * mg.hd;
    for (int idx = threadIdx.x; idx < n; idx += blockDim.x) {
        const double *row = soa + (long long)(idx >> 1) * g.pitch + GS;
        dst[idx] = side == 0 ? row[idx & 1] : row[g.nc - 2 + (idx & 1)];
    }
    __threadfence_system();
    __syncthreads();
    MailboxHeader *ph = (MailboxHeader *)peer, *mh = (MailboxHeader *)mine;
    __shared__ int ok;
    if (threadIdx.x == 0) {
        *(volatile unsigned long long *)&ph->halo_epoch[1 - side] = epoch;
        ok = wait_epoch(&mh->halo_epoch[side], epoch, mh) ? 1 : 0;
        __threadfence_system();
    }
    __syncthreads();
    if (!ok) return;
    const volatile double *src = (const volatile double *)(mine + mg.halo_off) + ((long long)side * 2 + par) * mg.hd;
    for (int idx = threadIdx.x; idx < n; idx += blockDim.x) {
        double *row = soa + (long long)(idx >> 1) * g.pitch + GS;
        if (side == 0) row[-2 + (idx & 1)] = src[idx];
        else row[g.nc + (idx & 1)] = src[idx];
    }
}

struct PeerTable {
    char *p[P2P_MAX_WORLD];
};

// all_reduce(MAX) of speed[nprob] over the ranks: one block; NaN propagates exactly as in the single-GPU reduction
__global__ void p2p_speed_max_kernel(MailboxGeom mg, PeerTable peers, int rank, double *speed, unsigned long long epoch)
{
    const int par = (int)(epoch & 1ull);
    const int total = mg.world * mg.nprob;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int r = i / mg.nprob, j = i - r * mg.nprob;
        double *slot = (double *)(peers.p[r] + mg.speed_off) + ((long long)par * mg.world + rank) * mg.nprob;
        slot[j] = speed[j];
    }
    __threadfence_system();
    __syncthreads();
    MailboxHeader *mh = (MailboxHeader *)peers.p[rank];
    __shared__ int bad;
    if (threadIdx.x == 0) bad = 0;
    __syncthreads();
    if (threadIdx.x < mg.world)
        *(volatile unsigned long long *)&((MailboxHeader *)peers.p[threadIdx.x])->speed_epoch[rank] = epoch;
    __syncthreads();  // every flag is out before any thread of this block starts to wait
    if (threadIdx.x < mg.world) {
        if (!wait_epoch(&mh->speed_epoch[threadIdx.x], epoch, mh)) bad = 1;
        __threadfence_system();
    }
    __syncthreads();
    if (bad) return;
    // maximum of the 64-bit patterns, as the single-GPU reduction takes it (atomicMax in the stage kernel): for the
    // non-negative speeds that is the numerical maximum, and a NaN (either sign) beats every finite value
    const volatile unsigned long long *all =
        (const volatile unsigned long long *)(peers.p[rank] + mg.speed_off) + (long long)par * mg.world * mg.nprob;
    for (int j = threadIdx.x; j < mg.nprob; j += blockDim.x) {
        unsigned long long m = all[j];
        for (int r = 1; r < mg.world; ++r) {
            const unsigned long long v = all[(long long)r * mg.nprob + j];
            m = (v > m) ? v : m;
        }
        ((unsigned long long *)speed)[j] = m;
    }
}

// The same reduction followed at once by the step control of the next step (step_control_one): one launch instead of two
// on the serial path between the last stage of a step and the first stage of the next.  `speed` enters as this rank's
// maximum; the control step uses the maximum over the ranks and clears `speed` for the next accumulation.
__global__ void p2p_speed_control_kernel(MailboxGeom mg, PeerTable peers, int rank, ControlArgs ca, unsigned long long epoch)
{
    const int par = (int)(epoch & 1ull);
    const int total = mg.world * mg.nprob;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int r = i / mg.nprob, j = i - r * mg.nprob;
        double *slot = (double *)(peers.p[r] + mg.speed_off) + ((long long)par * mg.world + rank) * mg.nprob;
        slot[j] = ca.speed[j];
    }
    __threadfence_system();
    __syncthreads();
    MailboxHeader *mh = (MailboxHeader *)peers.p[rank];
    __shared__ int bad;
    if (threadIdx.x == 0) bad = 0;
    __syncthreads();
    if (threadIdx.x < mg.world)
        *(volatile unsigned long long *)&((MailboxHeader *)peers.p[threadIdx.x])->speed_epoch[rank] = epoch;
    __syncthreads();
    if (threadIdx.x < mg.world) {
        if (!wait_epoch(&mh->speed_epoch[threadIdx.x], epoch, mh)) bad = 1;
        __threadfence_system();
    }
    __syncthreads();
    if (bad) return;
    const volatile unsigned long long *all =
        (const volatile unsigned long long *)(peers.p[rank] + mg.speed_off) + (long long)par * mg.world * mg.nprob;
    for (int j = threadIdx.x; j < mg.nprob; j += blockDim.x) {
        unsigned long long m = all[j];
        for (int r = 1; r < mg.world; ++r) {
            const unsigned long long v = all[(long long)r * mg.nprob + j];
            m = (v > m) ? v : m;
        }
        step_control_one(ca, j, __longlong_as_double((long long)m));
    }
}

struct dg1d_dist {
    ncclComm_t comm;
    int world, rank, left, right;  // neighbour ranks, -1 at a physical boundary
    cudaStream_t comm_stream;
    cudaEvent_t ev_edge, ev_int, ev_halo;
    const double *halo_valid;  // state buffer whose ghosts are up to date (nullptr: none)
    // peer-memory transport (dg1d_dist_p2p_open / _connect); NCCL is used until it is connected
    bool p2p;
    bool fused;  // peer transport: the halo exchange runs inside the stage kernel (DG1D_PEER_FUSED=1; default: its own kernel)
    MailboxGeom mg;
    char *mailbox;                  // mine (cudaMalloc)
    char *peer[P2P_MAX_WORLD];      // every rank's mailbox as mapped here; peer[rank] == mailbox
    unsigned long long halo_epoch, speed_epoch;
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
static int exchange(const dg1d_handle *h, dg1d_dist *d, double *soa, double *sl, double *sr, double *rl, double *rr,
                    bool on_caller_stream = false, cudaStream_t caller = nullptr)
{
    if (d->p2p) {
        cudaStream_t xs = on_caller_stream ? caller : d->comm_stream;  // (the caller's stream may be the null stream)
        HaloGeo g;
        g.nc = h->h.cfg.ncell; g.pitch = h->h.pitch; g.rows = h->h.nvar * h->h.nd * h->h.cfg.nprob;
        if (g.rows * 2 != d->mg.hd) { set_error("the peer mailbox was opened for another handle"); return DG1D_ESTATE; }
        d->halo_epoch += 1;
        p2p_halo_exchange_kernel<<<2, 128, 0, xs>>>(g, d->mg, soa, d->mailbox, d->left >= 0 ? d->peer[d->left] : nullptr,
                                                                d->right >= 0 ? d->peer[d->right] : nullptr, d->halo_epoch);
        CU_TRY(cudaGetLastError());
        count_launch();
        return DG1D_OK;
    }
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

static int reduce_speed(dg1d_dist *d, double *speed, int nprob, cudaStream_t st)
{
    if (d->p2p) {
        if (nprob != d->mg.nprob) { set_error("the peer mailbox was opened for %d problems, not %d", d->mg.nprob, nprob); return DG1D_ESTATE; }
        PeerTable pt;
        for (int r = 0; r < P2P_MAX_WORLD; ++r) pt.p[r] = r < d->world ? d->peer[r] : nullptr;
        d->speed_epoch += 1;
        p2p_speed_max_kernel<<<1, 128, 0, st>>>(d->mg, pt, d->rank, speed, d->speed_epoch);
        CU_TRY(cudaGetLastError());
        count_launch();
        return DG1D_OK;
    }
    NcclApi *n = nccl_api();
    // F9.  Reduced as 64-bit unsigned patterns: for non-negative doubles that order is the numerical one (max is exact),
    // and a NaN wave speed (a blown-up state, gas.py:23) wins on every rank -- ncclMax on doubles makes no promise about
    // NaN, and the run must stop at the same step whatever the transport and the rank that saw the NaN.
    NCCL_TRY(n->AllReduce(speed, speed, (size_t)nprob, ncclUint64, ncclMax, d->comm, st));
    return DG1D_OK;
}

static int dist_setup(dg1d_dist *d, const char id[128])
{
    NcclApi *n = nccl_api();
    ncclUniqueId u;
    memcpy(&u, id, 128);
    NCCL_TRY(n->CommInitRank(&d->comm, d->world, u, d->rank));
    int prio_lo = 0, prio_hi = 0;
    CU_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CU_TRY(cudaStreamCreateWithPriority(&d->comm_stream, cudaStreamNonBlocking, prio_hi));  // edge tiles + exchange go first
    CU_TRY(cudaEventCreateWithFlags(&d->ev_edge, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&d->ev_int, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&d->ev_halo, cudaEventDisableTiming));
    return DG1D_OK;
}

// peer transport: all-reduce(MAX) of the wave speed over the ranks and the step control of the next step, one kernel
static int speed_control(const dg1d_handle *h, dg1d_dist *d, const ControlArgs &ca, cudaStream_t st)
{
    if (h->h.cfg.nprob != d->mg.nprob) { set_error("the peer mailbox was opened for %d problems, not %d", d->mg.nprob, h->h.cfg.nprob); return DG1D_ESTATE; }
    PeerTable pt;
    for (int r = 0; r < P2P_MAX_WORLD; ++r) pt.p[r] = r < d->world ? d->peer[r] : nullptr;
    d->speed_epoch += 1;
    p2p_speed_control_kernel<<<1, 128, 0, st>>>(d->mg, pt, d->rank, ca, d->speed_epoch);
    CU_TRY(cudaGetLastError());
    count_launch();
    return DG1D_OK;
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
    if (!id || !out || world < 2 || rank < 0 || rank >= world || left < -1 || right < -1 || left >= world || right >= world) {
        set_error("bad argument (world >= 2, 0 <= rank < world, neighbours in -1 .. world-1)");
        return DG1D_EINVAL;
    }
    dg1d_dist *d = new (std::nothrow) dg1d_dist;
    if (!d) { set_error("out of host memory"); return DG1D_ENOMEM; }
    memset(d, 0, sizeof(*d));
    d->world = world; d->rank = rank; d->left = left; d->right = right;
    int rc = dist_setup(d, id);
    if (rc != DG1D_OK) {  // release whatever was built; the error text stays
        dg1d_dist_destroy(d);
        return rc;
    }
    *out = d;
    return DG1D_OK;
}

int dg1d_dist_p2p_open(dg1d_dist *d, const dg1d_handle *h, char ipc_handle[64])
{
    if (!d || !h || !ipc_handle) { set_error("null argument"); return DG1D_EINVAL; }
    if (d->world > P2P_MAX_WORLD) { set_error("peer transport supports at most %d ranks", P2P_MAX_WORLD); return DG1D_EINVAL; }
    if (d->mailbox) { set_error("the mailbox is already open"); return DG1D_ESTATE; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    d->mg = mailbox_geom((int)dg1d_halo_doubles(h), d->world, h->h.cfg.nprob);
    void *p = nullptr;
    CU_TRY(cudaMalloc(&p, (size_t)d->mg.bytes));  // a plain allocation: pool memory cannot be exported through CUDA IPC
    cudaIpcMemHandle_t hd;
    cudaError_t e = cudaMemset(p, 0, (size_t)d->mg.bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hd, p);
    if (e != cudaSuccess) {  // nothing is kept: a later call may try again
        set_error("opening the peer mailbox failed: %s", cudaGetErrorString(e));
        cudaFree(p);
        cudaGetLastError();
        return DG1D_ECUDA;
    }
    d->mailbox = (char *)p;
    memcpy(ipc_handle, &hd, 64);
    return DG1D_OK;
}

int dg1d_dist_p2p_connect(dg1d_dist *d, const char *ipc_handles)
{
    if (!d || !ipc_handles) { set_error("null argument"); return DG1D_EINVAL; }
    if (!d->mailbox) { set_error("dg1d_dist_p2p_open first"); return DG1D_ESTATE; }
    if (d->p2p) return DG1D_OK;
    for (int r = 0; r < d->world; ++r) {
        if (r == d->rank) { d->peer[r] = d->mailbox; continue; }
        cudaIpcMemHandle_t hd;
        memcpy(&hd, ipc_handles + 64 * (size_t)r, 64);
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            for (int q = 0; q < r; ++q)
                if (q != d->rank && d->peer[q]) { cudaIpcCloseMemHandle(d->peer[q]); d->peer[q] = nullptr; }
            set_error("cudaIpcOpenMemHandle for rank %d failed: %s (the NCCL transport stays in use)", r, cudaGetErrorString(e));
            return DG1D_ECUDA;
        }
        d->peer[r] = (char *)p;
    }
    d->p2p = true;
    const char *e = getenv("DG1D_PEER_FUSED");
    d->fused = e && atoi(e) != 0;
    return DG1D_OK;
}

int dg1d_dist_use_nccl(dg1d_dist *d)
{
    if (!d) { set_error("null argument"); return DG1D_EINVAL; }
    d->p2p = false;  // the mappings stay until dg1d_dist_destroy; the step loop goes back to ncclSend/Recv/AllReduce
    d->halo_valid = nullptr;
    return DG1D_OK;
}

int dg1d_dist_transport(const dg1d_dist *d) { return d && d->p2p ? DG1D_TRANSPORT_PEER : DG1D_TRANSPORT_NCCL; }

int dg1d_dist_p2p_error(dg1d_dist *d, int64_t *timed_out)
{
    if (!d || !timed_out) { set_error("null argument"); return DG1D_EINVAL; }
    *timed_out = 0;
    if (!d->mailbox) return DG1D_OK;
    // Both streams of the step loop are non-blocking streams, which a copy on the legacy default stream does not wait
    // for: drain the device first, so that a wait still pending in a queued exchange has either succeeded or timed out.
    CU_TRY(cudaDeviceSynchronize());
    unsigned long long e = 0;
    CU_TRY(cudaMemcpy(&e, d->mailbox + offsetof(MailboxHeader, error), sizeof(e), cudaMemcpyDeviceToHost));
    *timed_out = (int64_t)e;
    return DG1D_OK;
}

void dg1d_dist_destroy(dg1d_dist *d)
{
    if (!d) return;
    if (d->comm_stream) cudaStreamSynchronize(d->comm_stream);
    for (int r = 0; r < d->world && r < P2P_MAX_WORLD; ++r)
        if (r != d->rank && d->peer[r]) cudaIpcCloseMemHandle(d->peer[r]);
    if (d->mailbox) cudaFree(d->mailbox);
    if (nccl_api() && d->comm) nccl_api()->CommDestroy(d->comm);
    if (d->ev_edge) cudaEventDestroy(d->ev_edge);
    if (d->ev_int) cudaEventDestroy(d->ev_int);
    if (d->ev_halo) cudaEventDestroy(d->ev_halo);
    if (d->comm_stream) cudaStreamDestroy(d->comm_stream);
    cudaGetLastError();
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
    return reduce_speed(d, speed, nprob, (cudaStream_t)stream);
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
    if (d->p2p && d->fused) {
        // Peer transport, exchange inside the stage kernel (HaloLink, dg1d_kernels.cuh): one stream, the launches of a
        // single-GPU step plus the reduction.  Stage number e sends message e (its outermost cells of u_out) and first
        // waits for message e-1 (the ghosts of its u_in), unless those were just filled by a stand-alone exchange.
        if ((int)dg1d_halo_doubles(h) != d->mg.hd) { set_error("the peer mailbox was opened for another handle"); return DG1D_ESTATE; }
        static const bool fused_pdl = !(getenv("DG1D_DIST_PDL") && atoi(getenv("DG1D_DIST_PDL")) == 0);
        HaloLink link = {};
        link.mine = d->mailbox;
        link.peer_left = d->left >= 0 ? d->peer[d->left] : nullptr;
        link.peer_right = d->right >= 0 ? d->peer[d->right] : nullptr;
        link.halo_off = d->mg.halo_off;
        link.flag_off = (long long)offsetof(MailboxHeader, cell_epoch);
        link.err_off = (long long)offsetof(MailboxHeader, error);
        link.hd = d->mg.hd;
        for (int64_t s = 0; s < nsteps; ++s) {
            if (s == 0) {  // `speed` is the global maximum on entry
                if ((rc = dg1d_step_control(h, cfl, dx, Tf, max_steps, t, dt, lam, speed, it, done, dt_log, dt_cap, stream))) return rc;
            } else {       // reduction over the ranks + step control in one kernel
                const ControlArgs ca = {cfl, dx, Tf, (i64)max_steps, t, dt, lam, speed, (i64 *)it, done, dt_log, (i64)dt_cap};
                if ((rc = speed_control(h, d, ca, main))) return rc;
            }
            double *X = bufs[0], *Y = bufs[1], *Z = bufs[2];
            const double *uin[3] = {X, Y, Z}, *un[3] = {nullptr, X, X};
            double *uout[3] = {Y, Z, Y};
            for (int rk = 0; rk < 3; ++rk) {
                link.recv_epoch = d->halo_epoch;  // the last message sent carries the ghosts of this stage's input ...
                if (rk == 0 && d->halo_valid != X) {  // ... unless the state was replaced: exchange its ghosts first
                    if ((rc = exchange(h, d, X, send_left, send_right, recv_left, recv_right, true, main))) return rc;
                    link.recv_epoch = 0;
                }
                link.send_epoch = ++d->halo_epoch;
                StageArgs a = {uin[rk], un[rk], uout[rk], lam, done, (rk == 2) ? speed : nullptr, rk, MODE_STAGE, Mdx2, 0, fused_pdl, 0, link};
                if ((rc = launch_stage_any(h->h, a, main))) return rc;
            }
            bufs[0] = Y; bufs[1] = X; bufs[2] = Z;
            d->halo_valid = Y;
        }
        return reduce_speed(d, speed, h->h.cfg.nprob, main);  // as documented: `speed` is the global maximum on return
    }
    // Two streams.  `main`: the interior tiles of the three stages, back to back (each a programmatic dependent of the one
    // before).  `edge` (high priority): the edge tiles of a stage, then the exchange of their outermost cells, hidden behind
    // the interior.  The edges of stage rk are taken rk tiles deep (DG1D_PART_DEPTH): a tile reads two cells of each
    // neighbouring tile, so the interior of stage rk then reads only what the interior of stage rk-1 wrote, and `main`
    // waits for `edge` once per step (before the reduction, which needs every tile of stage 2) instead of at every stage.
    //   edges(rk)    needs the ghosts of its input (exchange(rk-1): same stream) and the interior tiles next to it
    //                (interior(rk-1): ev_int);
    //   interior(rk) needs interior(rk-1) only (same stream); interior(0) needs every tile of the previous step (ev_edge).
    static const bool deep = !(getenv("DG1D_DIST_DEEP_EDGES") && atoi(getenv("DG1D_DIST_DEEP_EDGES")) == 0);
    for (int64_t s = 0; s < nsteps; ++s) {
        if (d->p2p && s > 0) {  // `speed` holds this rank's maximum: reduce over the ranks and control the step, one kernel
            const ControlArgs ca = {cfl, dx, Tf, (i64)max_steps, t, dt, lam, speed, (i64 *)it, done, dt_log, (i64)dt_cap};
            if ((rc = speed_control(h, d, ca, main))) return rc;
        } else {                // `speed` is the global maximum (on entry; with NCCL: reduced at the end of the last step)
            if ((rc = dg1d_step_control(h, cfl, dx, Tf, max_steps, t, dt, lam, speed, it, done, dt_log, dt_cap, stream))) return rc;
        }
        CU_TRY(cudaEventRecord(d->ev_int, main));  // lam/done of this step; everything queued so far on main
        double *X = bufs[0], *Y = bufs[1], *Z = bufs[2];
        const double *uin[3] = {X, Y, Z}, *un[3] = {nullptr, X, X};
        double *uout[3] = {Y, Z, Y};
        for (int rk = 0; rk < 3; ++rk) {
            double *sp = (rk == 2) ? speed : nullptr;
            const int depth = deep ? DG1D_PART_DEPTH(rk) : 0;
            CU_TRY(cudaStreamWaitEvent(edge, d->ev_int, 0));
            if (rk == 0 && d->halo_valid != X) {  // first step, or the state was replaced: fill the ghosts of u^n
                if ((rc = exchange(h, d, X, send_left, send_right, recv_left, recv_right))) return rc;
            }
            if ((rc = dg1d_stage_part(h, DG1D_PART_EDGES | depth, rk, uin[rk], un[rk], uout[rk], lam, done, Mdx2, sp, edge))) return rc;
            CU_TRY(cudaEventRecord(d->ev_edge, edge));
            if ((rc = exchange(h, d, uout[rk], send_left, send_right, recv_left, recv_right))) return rc;
            if ((rc = dg1d_stage_part(h, DG1D_PART_INTERIOR | depth, rk, uin[rk], un[rk], uout[rk], lam, done, Mdx2, sp, stream))) return rc;
            if (rk < 2 || !deep) CU_TRY(cudaEventRecord(d->ev_int, main));
            if (rk == 2 || !deep) CU_TRY(cudaStreamWaitEvent(main, d->ev_edge, 0));  // the reduction / the next step needs the edge tiles
        }
        bufs[0] = Y; bufs[1] = X; bufs[2] = Z;
        d->halo_valid = Y;
        if (!d->p2p && (rc = reduce_speed(d, speed, h->h.cfg.nprob, main))) return rc;
    }
    if (d->p2p && (rc = reduce_speed(d, speed, h->h.cfg.nprob, main))) return rc;  // on return `speed` is the global maximum
    CU_TRY(cudaEventRecord(d->ev_halo, edge));
    CU_TRY(cudaStreamWaitEvent(main, d->ev_halo, 0));  // the caller's stream also sees the last ghosts
    return DG1D_OK;
}

}  // extern "C"
