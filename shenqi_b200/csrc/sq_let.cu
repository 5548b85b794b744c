// Multi-rank force tree: top-leaf moment exchange and the import of the reachable parts of remote sub-trees.
//
// Stands where the reference has
//   force_exchange_pseudodata / force_treeupdate_pseudos      libgadget/forcetree.cpp:1155-1206, 1212-1282   (row a7)
//   GravTopTreeWalk::toptree_visit + export_particle          libgadget/gravshort2.hpp:362-438, localtreewalk2.h:269-324 (a14)
//   TreeWalk::ev_toptree / ev_send_recv_export_import / ev_secondary / reduce   libgadget/treewalk2.h:618-812 (a15)
// Design (not the reference's): the reference exports every boundary particle to the owner of each pseudo node it
// opens, walks it there as a ghost and ships the partial result back - two all-to-alls per walk, with the remote walks
// on the critical path.  Here each rank imports ONCE PER TREE BUILD the part of every remote sub-tree its particles can
// reach (a locally essential tree): the owner marks, top-down and level by level, the nodes whose keep box
// (|centre - x| <= Rcut + len/2 on every axis: the reference's discard test, gravshort2.hpp:152-167, with the centre
// of mass inside the cell) meets the bounding box of a top leaf of the requesting rank, compacts them per requester
// (children stay contiguous, links rewritten to the compacted numbering), and sends nodes (cm, cl, lk: 80 B) and the
// particles of marked leaves (pos, mass: 28 B) with grouped ncclSend/ncclRecv straight into the tail of the
// receiver's node and particle arrays; the receiver shifts the links and hangs each imported root below its pseudo
// node.  A node that is NOT marked is discarded by every target of the requester, so leaving it out changes no
// interaction: the walk makes, per target, exactly the decisions of the reference's primary + ghost walks.
// NCCL is loaded with dlopen (the same libnccl.so.2 the process already has, e.g. torch's; else the system one), so
// the library itself links only cudart and cufft.
#include "sq_common.cuh"
#include <dlfcn.h>
#include <nccl.h>
#include <math.h>

// ------------------------------------------------------------------------------------ NCCL by dlopen
namespace {
struct nccl_api {
    void * handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char * (*GetErrorString)(ncclResult_t) = nullptr;
} NC;
}

static int sq_nccl_load()
{
    if(NC.handle) return SQ_OK;
    void * h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if(!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if(!h) { sq_set_error("multi-rank context: cannot load libnccl.so.2 (%s)", dlerror()); return SQ_ERR_ARG; }
#define SQ_NCSYM(name) do { *(void **) (&NC.name) = dlsym(h, "nccl" #name); if(!NC.name) { sq_set_error("libnccl: symbol nccl" #name " missing"); return SQ_ERR_ARG; } } while(0)
    SQ_NCSYM(GetUniqueId); SQ_NCSYM(CommInitRank); SQ_NCSYM(CommDestroy); SQ_NCSYM(AllReduce); SQ_NCSYM(AllGather);
    SQ_NCSYM(Send); SQ_NCSYM(Recv); SQ_NCSYM(GroupStart); SQ_NCSYM(GroupEnd); SQ_NCSYM(GetErrorString);
#undef SQ_NCSYM
    NC.handle = h;
    return SQ_OK;
}

#define SQ_NCCL(call)                                                                         \
    do {                                                                                      \
        ncclResult_t r_ = (call);                                                             \
        if(r_ != ncclSuccess) {                                                               \
            sq_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, NC.GetErrorString(r_)); \
            return SQ_ERR_CUDA;                                                               \
        }                                                                                     \
    } while(0)

// ------------------------------------------------------------------------------------ in-process loop-back transport
// TEST TRANSPORT: several ranks as host threads of ONE process (each with its own context, on one GPU or several) meet in
// a hub instead of NCCL; data moves by device-to-device copies.  It lets the whole multi-rank path (top nodes, moment
// exchange, marking, packing, import, walk) run under `pytest -m gpu` on a box with a single GPU, where NCCL refuses two
// ranks on one device.  Selected by an id made with sq_comm_loopback_id(); production runs use NCCL.
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
namespace {
struct sq_hub {
    int n = 0;
    std::mutex m;
    std::condition_variable cv;
    int arrived = 0;
    uint64_t gen = 0;
    std::vector<const void *> ptr;
    std::vector<std::vector<const void *>> sendp;
    void barrier()
    {
        std::unique_lock<std::mutex> lk(m);
        const uint64_t g = gen;
        if(++arrived == n) { arrived = 0; gen++; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g; });
    }
};
std::mutex hubs_lock;
std::map<uint64_t, std::shared_ptr<sq_hub>> hubs;
const char LOOP_MAGIC[8] = {'S', 'Q', 'L', 'O', 'O', 'P', 0, 1};
struct loop_ptrs { const void * p[32]; };
}

template <class T>
__global__ void let_loop_sum_kernel(loop_ptrs P, int n, size_t count, T * __restrict__ out)
{
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < count; i += (size_t) gridDim.x * blockDim.x) {
        T s = 0;
        for(int r = 0; r < n; r++) s += ((const T *) P.p[r])[i];
        out[i] = s;
    }
}

template <class T>
static int loop_allreduce(sq_ctx * c, T * dev, size_t count)
{
    sq_hub * H = (sq_hub *) c->loop_hub;
    sq_buf tmp;
    SQ_TRY(tmp.reserve(sizeof(T) * count));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    H->ptr[c->comm_rank] = dev;
    H->barrier();
    loop_ptrs P;
    for(int r = 0; r < H->n; r++) P.p[r] = H->ptr[r];
    let_loop_sum_kernel<T><<<c->num_sms * 2, 256, 0, c->stream>>>(P, H->n, count, tmp.as<T>());
    cudaError_t e = cudaStreamSynchronize(c->stream);
    H->barrier();
    if(e == cudaSuccess) e = cudaMemcpyAsync(dev, tmp.p, sizeof(T) * count, cudaMemcpyDeviceToDevice, c->stream);
    if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    tmp.release();
    if(e != cudaSuccess) { sq_set_error("loop-back all-reduce: %s", cudaGetErrorString(e)); return SQ_ERR_CUDA; }
    return SQ_OK;
}

int sq_comm_allreduce_sum_i64(sq_ctx * c, int64_t * dev, size_t n)
{
    if(c->nranks <= 1 || n == 0) return SQ_OK;
    if(c->loop_hub) return loop_allreduce<long long>(c, (long long *) dev, n);
    SQ_NCCL(NC.AllReduce(dev, dev, n, ncclInt64, ncclSum, (ncclComm_t) c->nccl_comm, c->stream));
    return SQ_OK;
}
int sq_comm_allreduce_sum_f64(sq_ctx * c, double * dev, size_t n)
{
    if(c->nranks <= 1 || n == 0) return SQ_OK;
    if(c->loop_hub) return loop_allreduce<double>(c, dev, n);
    SQ_NCCL(NC.AllReduce(dev, dev, n, ncclDouble, ncclSum, (ncclComm_t) c->nccl_comm, c->stream));
    return SQ_OK;
}
// dst[r * count .. ] = src of rank r
static int sq_comm_allgather_i64(sq_ctx * c, const int64_t * src, int64_t * dst, size_t count)
{
    if(c->loop_hub) {
        sq_hub * H = (sq_hub *) c->loop_hub;
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        H->ptr[c->comm_rank] = src;
        H->barrier();
        cudaError_t e = cudaSuccess;
        for(int r = 0; r < H->n && e == cudaSuccess; r++)
            e = cudaMemcpyAsync(dst + (size_t) r * count, H->ptr[r], sizeof(int64_t) * count, cudaMemcpyDefault, c->stream);
        if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        H->barrier();
        if(e != cudaSuccess) { sq_set_error("loop-back all-gather: %s", cudaGetErrorString(e)); return SQ_ERR_CUDA; }
        return SQ_OK;
    }
    SQ_NCCL(NC.AllGather(src, dst, count, ncclInt64, (ncclComm_t) c->nccl_comm, c->stream));
    return SQ_OK;
}
// every rank sends sbytes[r] bytes at sptr[r] to rank r and receives rbytes[r] bytes from it at rptr[r]
static int sq_comm_exchange(sq_ctx * c, const std::vector<const unsigned char *> & sptr, const std::vector<size_t> & sbytes,
                            const std::vector<unsigned char *> & rptr, const std::vector<size_t> & rbytes)
{
    const int R = c->nranks, me = c->comm_rank;
    if(c->loop_hub) {
        sq_hub * H = (sq_hub *) c->loop_hub;
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        for(int r = 0; r < R; r++) H->sendp[me][r] = sptr[r];
        H->barrier();
        cudaError_t e = cudaSuccess;
        for(int s = 0; s < R && e == cudaSuccess; s++)
            if(s != me && rbytes[s] > 0) e = cudaMemcpyAsync(rptr[s], H->sendp[s][me], rbytes[s], cudaMemcpyDefault, c->stream);
        if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        H->barrier();
        if(e != cudaSuccess) { sq_set_error("loop-back exchange: %s", cudaGetErrorString(e)); return SQ_ERR_CUDA; }
        return SQ_OK;
    }
    SQ_NCCL(NC.GroupStart());
    for(int r = 0; r < R; r++) {
        if(r == me) continue;
        if(sbytes[r] > 0) SQ_NCCL(NC.Send(sptr[r], sbytes[r], ncclUint8, r, (ncclComm_t) c->nccl_comm, c->stream));
        if(rbytes[r] > 0) SQ_NCCL(NC.Recv(rptr[r], rbytes[r], ncclUint8, r, (ncclComm_t) c->nccl_comm, c->stream));
    }
    SQ_NCCL(NC.GroupEnd());
    return SQ_OK;
}
void sq_comm_destroy(sq_ctx * c)
{
    if(c->nccl_comm && NC.CommDestroy) NC.CommDestroy((ncclComm_t) c->nccl_comm);
    c->nccl_comm = nullptr;
    c->loop_hub = nullptr;
}

/* TEST TRANSPORT id: ranks of one process (host threads) that pass the same id meet in one hub; `key` tells hubs apart. */
extern "C" int sq_comm_loopback_id(void * id128, int key)
{
    if(!id128) return SQ_ERR_ARG;
    memset(id128, 0, SQ_NCCL_UNIQUE_ID_BYTES);
    memcpy(id128, LOOP_MAGIC, 8);
    const uint64_t k = (uint64_t) (unsigned) key;
    memcpy((char *) id128 + 8, &k, 8);
    return SQ_OK;
}

static_assert(sizeof(ncclUniqueId) == SQ_NCCL_UNIQUE_ID_BYTES, "ncclUniqueId is 128 bytes");

extern "C" int sq_comm_unique_id(void * id128)
{
    if(!id128) return SQ_ERR_ARG;
    SQ_TRY(sq_nccl_load());
    ncclUniqueId id;
    SQ_NCCL(NC.GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return SQ_OK;
}

extern "C" int sq_ctx_create_dist(sq_ctx ** out, int device, const void * nccl_unique_id, int rank, int nranks)
{
    if(!out || nranks < 1 || rank < 0 || rank >= nranks || nranks > 32) { sq_set_error("sq_ctx_create_dist: rank %d of %d (at most 32 ranks)", rank, nranks); return SQ_ERR_ARG; }
    SQ_TRY(sq_ctx_create(out, device));
    sq_ctx * c = *out;
    c->comm_rank = rank; c->nranks = nranks;
    if(nranks > 1) {
        if(!nccl_unique_id) { sq_set_error("sq_ctx_create_dist: null ncclUniqueId"); sq_ctx_destroy(c); *out = nullptr; return SQ_ERR_ARG; }
        if(memcmp(nccl_unique_id, LOOP_MAGIC, 8) == 0) {
            uint64_t key;
            memcpy(&key, (const char *) nccl_unique_id + 8, 8);
            std::lock_guard<std::mutex> g(hubs_lock);
            std::shared_ptr<sq_hub> & h = hubs[key];
            if(!h || h->n != nranks) {
                h = std::make_shared<sq_hub>();
                h->n = nranks; h->ptr.assign(nranks, nullptr); h->sendp.assign(nranks, std::vector<const void *>(nranks, nullptr));
            }
            c->loop_hub = h.get();
            return SQ_OK;
        }
        int rc = sq_nccl_load();
        if(rc == SQ_OK) {
            ncclUniqueId id;
            memcpy(&id, nccl_unique_id, sizeof(id));
            ncclComm_t comm = nullptr;
            ncclResult_t r = NC.CommInitRank(&comm, nranks, id, rank);
            if(r != ncclSuccess) { sq_set_error("ncclCommInitRank(rank %d of %d): %s", rank, nranks, NC.GetErrorString(r)); rc = SQ_ERR_CUDA; }
            else c->nccl_comm = comm;
        }
        if(rc != SQ_OK) { sq_ctx_destroy(c); *out = nullptr; return rc; }
    }
    return SQ_OK;
}

extern "C" int sq_comm_rank(sq_ctx * c, int * rank, int * nranks)
{
    if(!c) return SQ_ERR_ARG;
    if(rank) *rank = c->comm_rank;
    if(nranks) *nranks = c->nranks;
    return SQ_OK;
}

// ------------------------------------------------------------------------------------ kernels
static int let_grid(const sq_ctx * c, long long n, int per_sm = 8)
{
    long long need = (n + 255) / 256, cap = (long long) c->num_sms * per_sm;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}

__host__ __device__ __forceinline__ unsigned let_quant(double x, double qinv)
{
    return (unsigned) (long long) floor(x * qinv);
}

// per top leaf of this rank: moments of its node (force_exchange_pseudodata, forcetree.cpp:1159-1172) and the bounding box
// of its particles in 32-bit fixed point (box / 2^32); leaves of other ranks stay zero so that a sum over ranks gathers
struct let_leafdata { double cm[4]; long long lo[3], hi[3]; long long valid, pad; };     // 12 x 8 B
static_assert(sizeof(let_leafdata) == 96, "leaf record is 12 words");

// bounding boxes: every particle finds its top leaf (the leaves' particle ranges are consecutive in the sorted order),
// a warp reduces over the lanes that share the first lane's leaf, one set of atomics per warp and leaf
__global__ void __launch_bounds__(256) let_leafbox_kernel(const int2 * __restrict__ leafrange, const int * __restrict__ myleaves, int nmine,
                                                          const double * __restrict__ pos, long long n, double qinv, unsigned * __restrict__ box /* [leaf][6] */)
{
    for(long long i0 = (blockIdx.x * (long long) blockDim.x + threadIdx.x) & ~31ll; i0 < n; i0 += (long long) gridDim.x * blockDim.x) {
        const long long i = i0 + sq_lane();
        int leaf = -1;
        unsigned q[3] = {0, 0, 0};
        if(i < n) {
            // last of my leaves (sorted by first particle) whose range starts at or before i
            int lo = 0, hi = nmine;
            while(hi - lo > 1) { const int mid = (lo + hi) >> 1; if(leafrange[myleaves[mid]].x <= i) lo = mid; else hi = mid; }
            leaf = myleaves[lo];
            const int2 r = leafrange[leaf];
            if(i < r.x || i >= (long long) r.x + r.y) leaf = -1;
            #pragma unroll
            for(int a = 0; a < 3; a++) q[a] = let_quant(pos[3 * i + a], qinv);
        }
        unsigned todo = __ballot_sync(SQ_FULL, leaf >= 0);
        while(todo) {
            const int lead = __ffs((int) todo) - 1;
            const int l0 = __shfl_sync(SQ_FULL, leaf, lead);
            const bool in = leaf == l0;
            const unsigned grp = __ballot_sync(SQ_FULL, in);
            #pragma unroll
            for(int a = 0; a < 3; a++) {
                const unsigned mn = __reduce_min_sync(SQ_FULL, in ? q[a] : 0xffffffffu), mx = __reduce_max_sync(SQ_FULL, in ? q[a] : 0u);
                if((int) sq_lane() == lead) { atomicMin(&box[6 * l0 + a], mn); atomicMax(&box[6 * l0 + 3 + a], mx); }
            }
            todo &= ~grp;
        }
    }
}

__global__ void let_leafdata_kernel(const sq_topleaf * __restrict__ tl, int nleaves, int myrank, const int * __restrict__ treenode,
                                    const int2 * __restrict__ leafrange, const double4 * __restrict__ cm, const unsigned * __restrict__ box,
                                    let_leafdata * __restrict__ out)
{
    for(int leaf = blockIdx.x * blockDim.x + threadIdx.x; leaf < nleaves; leaf += gridDim.x * blockDim.x) {
        let_leafdata d;
        memset(&d, 0, sizeof(d));
        if(tl[leaf].Task == myrank) {
            const double4 m = cm[treenode[leaf]];
            d.cm[0] = m.x; d.cm[1] = m.y; d.cm[2] = m.z; d.cm[3] = m.w;
            if(leafrange[leaf].y > 0) {
                for(int a = 0; a < 3; a++) { d.lo[a] = box[6 * leaf + a]; d.hi[a] = box[6 * leaf + 3 + a]; }
                d.valid = 1;
            }
        }
        out[leaf] = d;
    }
}

// moments of the other ranks' top leaves into their pseudo nodes (forcetree.cpp:1190-1203)
__global__ void let_scatter_moments_kernel(const sq_topleaf * __restrict__ tl, int nleaves, int myrank, const int * __restrict__ treenode,
                                           const let_leafdata * __restrict__ ld, double4 * __restrict__ cm)
{
    for(int leaf = blockIdx.x * blockDim.x + threadIdx.x; leaf < nleaves; leaf += gridDim.x * blockDim.x)
        if(tl[leaf].Task != myrank)
            cm[treenode[leaf]] = make_double4(ld[leaf].cm[0], ld[leaf].cm[1], ld[leaf].cm[2], ld[leaf].cm[3]);
}

// One level of the top-down marking: mask[i] bit r = rank r may reach node i (its keep box meets the particle box of a
// top leaf of r).  Only nodes of this rank's own sub-trees are marked.  e_level = keep half-width in quanta (+ slack),
// negative = every target keeps (Rcut + len/2 >= Box/2).
struct let_mark_par { long long e_level[SQ_MAXLEVEL + 3]; unsigned others; int myrank, nranks; };

__device__ __forceinline__ bool let_axis_near(unsigned c, unsigned lo, unsigned hi, unsigned long long e)
{
    if((unsigned) (c - lo) <= (unsigned) (hi - lo)) return true;            // inside [lo, hi]
    const unsigned up = lo - c, down = c - hi;                                 // periodic distances to the two ends
    return (unsigned long long) (up < down ? up : down) <= e;
}

__global__ void __launch_bounds__(256) let_mark_kernel(const int4 * __restrict__ lk, const double4 * __restrict__ cl, const int * __restrict__ parent,
                                                       int ls, int le, double qinv, const let_leafdata * __restrict__ ld, const sq_task * __restrict__ tasks,
                                                       unsigned * __restrict__ mask, const let_mark_par P)
{
    for(int i = ls + blockIdx.x * blockDim.x + threadIdx.x; i < le; i += gridDim.x * blockDim.x) {
        const int4 L = lk[i];
        unsigned base, m = 0;
        if(SQ_INFO_TOP(L.w)) base = (!SQ_INFO_ITOP(L.w) && SQ_INFO_TYPE(L.w) != 2) ? P.others : 0u;   // a top leaf of this rank starts a sub-tree
        else base = mask[parent[i]];
        if(base) {
            const long long e = P.e_level[SQ_INFO_LEVEL(L.w)];
            if(e < 0) m = base;
            else {
                const double4 C = cl[i];
                const unsigned cx = let_quant(C.x, qinv), cy = let_quant(C.y, qinv), cz = let_quant(C.z, qinv);
                unsigned todo = base;
                while(todo) {
                    const int r = __ffs((int) todo) - 1;
                    todo &= todo - 1u;
                    for(int leaf = tasks[r].StartLeaf; leaf < tasks[r].EndLeaf; leaf++) {
                        const let_leafdata & B = ld[leaf];
                        if(!B.valid) continue;
                        if(let_axis_near(cx, (unsigned) B.lo[0], (unsigned) B.hi[0], (unsigned long long) e) &&
                           let_axis_near(cy, (unsigned) B.lo[1], (unsigned) B.hi[1], (unsigned long long) e) &&
                           let_axis_near(cz, (unsigned) B.lo[2], (unsigned) B.hi[2], (unsigned long long) e)) { m |= 1u << r; break; }
                    }
                }
            }
        }
        mask[i] = m;
    }
}

// flags for one requester: node flag and particle count of marked leaves
__global__ void let_flags_kernel(const unsigned * __restrict__ mask, const int4 * __restrict__ lk, int n, int r, uint32_t * __restrict__ fn, uint32_t * __restrict__ fp)
{
    for(int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const bool f = (mask[i] >> r) & 1u;
        fn[i] = f ? 1u : 0u;
        const int4 L = lk[i];
        fp[i] = (f && SQ_INFO_TYPE(L.w) == 0) ? (uint32_t) L.y : 0u;
    }
}

// pack the marked nodes (compacted numbering: children of a node stay contiguous and in octant order) and the particles
// of marked leaves; sn / sp = exclusive scans of the flags (one extra entry = totals)
__global__ void __launch_bounds__(256) let_pack_kernel(const unsigned * __restrict__ mask, int r, const int4 * __restrict__ lk, const double4 * __restrict__ cm,
                                                       const double4 * __restrict__ cl, const double * __restrict__ pos, const float * __restrict__ mass, int n,
                                                       const uint32_t * __restrict__ sn, const uint32_t * __restrict__ sp, double4 * __restrict__ ocm,
                                                       double4 * __restrict__ ocl, int4 * __restrict__ olk, double * __restrict__ opos, float * __restrict__ omass)
{
    for(int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if(!((mask[i] >> r) & 1u)) continue;
        const uint32_t j = sn[i];
        const int4 L = lk[i];
        ocm[j] = cm[i]; ocl[j] = cl[i];
        int4 O = make_int4(0, 0, -1, L.w & 0x00ffffff);
        if(SQ_INFO_TYPE(L.w) == 1) {
            unsigned cmk = 0;
            for(int k = 0; k < L.y; k++)
                if((mask[L.x + k] >> r) & 1u) cmk |= 1u << SQ_INFO_OCT(lk[L.x + k].w);
            O.x = (int) sn[L.x];
            O.y = (int) (sn[L.x + L.y] - sn[L.x]);
            O.w |= (int) (cmk << 24);
        } else {
            const uint32_t q = sp[i];
            O.x = (int) q; O.y = L.y;
            for(int k = 0; k < L.y; k++) {
                const long long p = L.x + k;
                opos[3 * (long long) (q + k)] = pos[3 * p]; opos[3 * (long long) (q + k) + 1] = pos[3 * p + 1]; opos[3 * (long long) (q + k) + 2] = pos[3 * p + 2];
                omass[q + k] = mass[p];
            }
        }
        olk[j] = O;
    }
}

// compacted number of each of this rank's top-leaf nodes for requester r, or -1
__global__ void let_roots_kernel(const sq_topleaf * __restrict__ tl, int nleaves, int myrank, const int * __restrict__ treenode, const unsigned * __restrict__ mask,
                                 int r, const uint32_t * __restrict__ sn, int * __restrict__ rootidx)
{
    for(int leaf = blockIdx.x * blockDim.x + threadIdx.x; leaf < nleaves; leaf += gridDim.x * blockDim.x) {
        int v = -1;
        if(tl[leaf].Task == myrank) {
            const int no = treenode[leaf];
            if((mask[no] >> r) & 1u) v = (int) sn[no];
        }
        rootidx[leaf] = v;
    }
}

// receiver: shift the links of an imported block to its place in the combined arrays
__global__ void let_relocate_kernel(int4 * __restrict__ lk, long long noff, long long nn, long long poff)
{
    for(long long j = blockIdx.x * (long long) blockDim.x + threadIdx.x; j < nn; j += (long long) gridDim.x * blockDim.x) {
        int4 L = lk[noff + j];
        L.x += (int) (SQ_INFO_TYPE(L.w) == 1 ? noff : poff);
        lk[noff + j] = L;
    }
}
// receiver: the imported root of a top leaf replaces the pseudo node's links (moments and cell stay: they are the owner's)
__global__ void let_attach_kernel(const sq_task * __restrict__ tasks, int src, const int * __restrict__ rootidx, const int * __restrict__ treenode,
                                  int4 * __restrict__ lk, long long noff)
{
    for(int leaf = tasks[src].StartLeaf + blockIdx.x * blockDim.x + threadIdx.x; leaf < tasks[src].EndLeaf; leaf += gridDim.x * blockDim.x) {
        const int ri = rootidx[leaf];
        if(ri < 0) continue;
        const int pn = treenode[leaf];
        const int4 S = lk[noff + ri], D = lk[pn];
        // type, child mask and level from the import; TopLevel flag and sibling of the pseudo node kept
        lk[pn] = make_int4(S.x, S.y, D.z, (S.w & ~(3 << 8) & ~(1 << 12)) | (SQ_INFO_TYPE(S.w) << 8) | (1 << 12));
    }
}

// ------------------------------------------------------------------------------------ toptree export count (row a14)
// GravTopTreeWalk::toptree_visit<TOPTREE_COUNT>, gravshort2.hpp:362-438, one thread per local particle over the top-level
// nodes (FP64, the reference's own expressions): counts the pseudo nodes a particle would open = its exports.
struct let_top_par { double box, half, rcut, rcut2, bh2, errtol; int relative, myrank; };
__device__ __forceinline__ double let_nearest(double x, double box, double half) { return (x > half) ? (x - box) : ((x < -half) ? (x + box) : x); }

__global__ void __launch_bounds__(256) let_toptree_count_kernel(const int4 * __restrict__ lk, const double4 * __restrict__ cm, const double4 * __restrict__ cl,
                                                                const int * __restrict__ topidx, const sq_topnode * __restrict__ tn, const sq_topleaf * __restrict__ tl,
                                                                const double * __restrict__ pos, const double * __restrict__ oldacc, long long n,
                                                                unsigned long long * __restrict__ out, const let_top_par P)
{
    unsigned long long nexp = 0, npart = 0;
    for(long long i = blockIdx.x * (long long) blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const double x = pos[3 * i], y = pos[3 * i + 1], z = pos[3 * i + 2], aold = P.errtol * oldacc[i];
        unsigned tasks_seen = 0;
        int no = 0;
        while(no >= 0) {
            const int4 L = lk[no];
            const double4 M = cm[no], C = cl[no];
            const double dx = let_nearest(M.x - x, P.box, P.half), dy = let_nearest(M.y - y, P.box, P.half), dz = let_nearest(M.z - z, P.box, P.half);
            const double r2 = dx * dx + dy * dy + dz * dz;
            const double ex = fabs(let_nearest(C.x - x, P.box, P.half)), ey = fabs(let_nearest(C.y - y, P.box, P.half)), ez = fabs(let_nearest(C.z - z, P.box, P.half));
            bool open = true;
            if(r2 > P.rcut2) { const double eff = P.rcut + 0.5 * C.w; if(ex > eff || ey > eff || ez > eff) open = false; }   // discarded
            if(open) {
                const double len = C.w, inside = 0.6 * len;
                open = (P.relative && (M.w * len * len > r2 * r2 * aold)) || (len * len / r2 > P.bh2) || (ex < inside && ey < inside && ez < inside);
            }
            if(!open) { no = L.z; continue; }
            const int ti = topidx[no];
            const sq_topnode T = tn[ti >= 0 ? ti : 0];
            if(ti >= 0 && T.Daughter < 0) {
                // a top leaf: a pseudo node that would be opened is an export (coalesced per task, localtreewalk2.h:269-324)
                const int task = tl[T.Leaf].Task;
                if(task != P.myrank) tasks_seen |= 1u << task;
                no = L.z;
                continue;
            }
            no = L.x;       // internal top node: its 8 children are top nodes
        }
        nexp += __popc(tasks_seen);
        npart += tasks_seen ? 1 : 0;
    }
    for(int o = 16; o >= 1; o >>= 1) { nexp += __shfl_xor_sync(SQ_FULL, nexp, o); npart += __shfl_xor_sync(SQ_FULL, npart, o); }
    if(sq_lane() == 0) { atomicAdd(&out[0], nexp); atomicAdd(&out[1], npart); }
}

// ------------------------------------------------------------------------------------ host
static double let_ms(cudaEvent_t a, cudaEvent_t b) { float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }

static void let_keep_levels(const sq_ctx * c, let_mark_par & P, double qinv)
{
    // Rcut as the walk will use it (GravTreeParams, gravshort2.hpp:45-47); + 16 quanta of slack for the rounding of centres and boxes
    const sq_params & par = c->par;
    const double rcut = c->have_raw ? c->raw.Rcut : par.TreeRcut * par.Asmth * (par.BoxSize / par.Nmesh);
    for(int L = 0; L < SQ_MAXLEVEL + 3; L++) {
        const double len = ldexp(1.001 * par.BoxSize, -L);
        const double e = (rcut + 0.5 * len) * (1.0 + 1e-12) * qinv + 16.0;
        P.e_level[L] = e >= 2147483648.0 ? -1 : (long long) ceil(e);
    }
}

extern "C" int sq_tree_build_dist(sq_ctx * c, sq_tree_info * info, sq_let_info * let)
{
    if(!c) return SQ_ERR_ARG;
    if(!c->have_domain) { sq_set_error("sq_tree_build_dist: install a domain first (sq_domain_set)"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaSetDevice(c->device));
    sq_let_info LI;
    memset(&LI, 0, sizeof(LI));
    cudaEvent_t * ev = c->ev;
    SQ_CUDA(cudaEventRecord(ev[10], c->stream));
    SQ_TRY(sq_tree_run(c, nullptr, 0, info));
    SQ_CUDA(cudaEventRecord(ev[11], c->stream));
    const int NL = (int) c->dom_leaves.size(), R = c->nranks, me = c->comm_rank;
    // particles in tree order: the sorted arrays, or their geometric re-ordering when some particle lies on a cell face
    sq_buf & SP = c->tree_identity ? c->pos : c->tree_pos;
    sq_buf & SM = c->tree_identity ? c->mass : c->tree_mass;
    const int64_t n = c->tree_np, nn0 = c->tree_numnodes;
    LI.local_nodes = nn0; LI.local_particles = n;
    const double qinv = 4294967296.0 / c->par.BoxSize;
    sq_buf & d_tasks = c->let_boxes;          // task table on the device
    SQ_TRY(d_tasks.reserve(sizeof(sq_task) * (size_t) R));
    SQ_CUDA(cudaMemcpyAsync(d_tasks.p, c->dom_tasks.data(), sizeof(sq_task) * (size_t) R, cudaMemcpyHostToDevice, c->stream));

    // ---- a7: top-leaf moments (and particle boxes) of every rank; internal top nodes recomputed
    SQ_TRY(c->let_leafdata.reserve(sizeof(let_leafdata) * (size_t) (NL + 1)));
    let_leafdata * ld = c->let_leafdata.as<let_leafdata>();
    {
        // my leaves in the order of their particle ranges (= leaf order: the domain's leaves are in key order)
        std::vector<int> mine;
        for(int l = c->dom_tasks[me].StartLeaf; l < c->dom_tasks[me].EndLeaf; l++) mine.push_back(l);
        sq_buf & scratch = c->let_scan_p;
        SQ_TRY(scratch.reserve(sizeof(unsigned) * 6 * (size_t) NL + sizeof(int) * (size_t) (mine.size() + 1) + 64));
        unsigned * d_box = scratch.as<unsigned>();
        int * d_mine = (int *) (d_box + 6 * (size_t) NL);
        std::vector<unsigned> init(6 * (size_t) NL);
        for(int l = 0; l < NL; l++) for(int a = 0; a < 3; a++) { init[6 * l + a] = 0xffffffffu; init[6 * l + 3 + a] = 0u; }
        SQ_CUDA(cudaMemcpyAsync(d_box, init.data(), sizeof(unsigned) * init.size(), cudaMemcpyHostToDevice, c->stream));
        SQ_CUDA(cudaMemcpyAsync(d_mine, mine.data(), sizeof(int) * mine.size(), cudaMemcpyHostToDevice, c->stream));
        if(n > 0 && !mine.empty()) {
            let_leafbox_kernel<<<let_grid(c, n), 256, 0, c->stream>>>(c->d_leafrange.as<int2>(), d_mine, (int) mine.size(), SP.as<double>(), (long long) n, qinv, d_box);
            c->kernel_launches++;
        }
        let_leafdata_kernel<<<let_grid(c, NL), 256, 0, c->stream>>>(c->d_topleaves.as<sq_topleaf>(), NL, me, c->d_treenode.as<int>(), c->d_leafrange.as<int2>(),
                c->t_cm.as<double4>(), d_box, ld);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    if(R > 1) {
        // each record has exactly one non-zero contributor (its owner): the sum over ranks is a gather, bit for bit
        // (doubles as 64-bit integers: integer addition of zeros leaves every bit pattern alone)
        SQ_TRY(sq_comm_allreduce_sum_i64(c, (int64_t *) ld, (size_t) NL * (sizeof(let_leafdata) / 8)));
        let_scatter_moments_kernel<<<let_grid(c, NL), 256, 0, c->stream>>>(c->d_topleaves.as<sq_topleaf>(), NL, me, c->d_treenode.as<int>(), ld, c->t_cm.as<double4>());
        c->kernel_launches++;
        for(int L = c->dom_maxlevel - 1; L >= 0; L--) {
            const int ls = c->lvl_start[L], le = c->lvl_start[L + 1];
            // top nodes of level L are the first nodes of the level only if the level holds nothing else; scan the whole level
            sq_tree_moments_top(c, ls, le);
        }
        SQ_CUDA(cudaGetLastError());
    }
    // mass unit of the walk's FP32 tiers: mean particle mass of the whole box
    {
        double4 h_root = make_double4(0, 0, 0, 0);
        int64_t * d_n = (int64_t *) (ld + NL);
        const int64_t h_n = n;
        SQ_CUDA(cudaMemcpyAsync(d_n, &h_n, 8, cudaMemcpyHostToDevice, c->stream));
        SQ_TRY(sq_comm_allreduce_sum_i64(c, d_n, 1));
        int64_t ntot = 0;
        SQ_CUDA(cudaMemcpyAsync(&ntot, d_n, 8, cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaMemcpyAsync(&h_root, c->t_cm.p, sizeof(h_root), cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        c->tree_total_mass = h_root.w;
        c->walk_munit = (ntot > 0 && h_root.w > 0) ? h_root.w / (double) ntot : 0.0;
    }
    SQ_CUDA(cudaEventRecord(ev[12], c->stream));
    c->n_local_nodes = nn0; c->n_sources = n; c->let_attached = false;
    if(R == 1) {
        if(let) { LI.build_ms = let_ms(ev[10], ev[11]); *let = LI; }
        return SQ_OK;
    }

    // ---- a14 / a15 (import form): mark what each other rank can reach
    SQ_TRY(c->let_mask.reserve(sizeof(unsigned) * (size_t) (nn0 + 1)));
    let_mark_par MP;
    memset(&MP, 0, sizeof(MP));
    let_keep_levels(c, MP, qinv);
    MP.myrank = me; MP.nranks = R;
    MP.others = (R >= 32 ? 0xffffffffu : ((1u << R) - 1u)) & ~(1u << me);
    for(int L = 0; L <= c->tree_maxlevel; L++) {
        const int ls = c->lvl_start[L], le = c->lvl_start[L + 1];
        if(le <= ls) continue;
        let_mark_kernel<<<let_grid(c, le - ls), 256, 0, c->stream>>>(c->t_lk.as<int4>(), c->t_cl.as<double4>(), c->t_parent.as<int>(), ls, le, qinv, ld,
                d_tasks.as<sq_task>(), c->let_mask.as<unsigned>(), MP);
        c->kernel_launches++;
    }
    SQ_CUDA(cudaGetLastError());
    SQ_CUDA(cudaEventRecord(ev[13], c->stream));

    // ---- per requester: compact and pack.  Send buffer = [cm | cl | lk | pos | mass | rootidx] per peer
    SQ_TRY(c->let_scan_n.reserve(sizeof(uint32_t) * (size_t) (nn0 + 2)));
    SQ_TRY(c->let_scan_p.reserve(sizeof(uint32_t) * (size_t) (nn0 + 2)));
    std::vector<int64_t> snn(R, 0), snp(R, 0), soff(R, 0);
    // every section starts 32-byte aligned, every peer block is a multiple of 256 bytes
    auto pad32 = [](size_t x) { return (x + 31) & ~(size_t) 31; };
    auto peer_bytes = [NL, pad32](int64_t a, int64_t b) { return (((size_t) a * 80 + pad32((size_t) b * 24) + pad32((size_t) b * 4) + pad32((size_t) NL * 4)) + 255) & ~(size_t) 255; };
    // pass 1: sizes
    for(int r = 0; r < R; r++) {
        if(r == me) continue;
        uint32_t * fn = c->let_scan_n.as<uint32_t>(), * fp = c->let_scan_p.as<uint32_t>();
        let_flags_kernel<<<let_grid(c, nn0), 256, 0, c->stream>>>(c->let_mask.as<unsigned>(), c->t_lk.as<int4>(), (int) nn0, r, fn, fp);
        c->kernel_launches++;
        SQ_TRY(sq_exclusive_scan_u32(c, fn, nn0, fn + nn0));
        SQ_TRY(sq_exclusive_scan_u32(c, fp, nn0, fp + nn0));
        uint32_t h[2] = {0, 0};
        SQ_CUDA(cudaMemcpyAsync(&h[0], fn + nn0, 4, cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaMemcpyAsync(&h[1], fp + nn0, 4, cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        snn[r] = h[0]; snp[r] = h[1];
    }
    size_t send_total = 0;
    for(int r = 0; r < R; r++) { soff[r] = (int64_t) send_total; if(r != me) send_total += peer_bytes(snn[r], snp[r]); }
    SQ_TRY(c->let_send.reserve(send_total + 256));
    auto sect = [&](unsigned char * base, int64_t a, int64_t b, double4 *& pcm, double4 *& pcl, int4 *& plk, double *& ppos, float *& pmass, int *& proot) {
        unsigned char * q = base;
        pcm = (double4 *) q; q += (size_t) a * 32;
        pcl = (double4 *) q; q += (size_t) a * 32;
        plk = (int4 *) q; q += (size_t) a * 16;
        ppos = (double *) q; q += pad32((size_t) b * 24);
        pmass = (float *) q; q += pad32((size_t) b * 4);
        proot = (int *) q;
    };
    // pass 2: pack
    for(int r = 0; r < R; r++) {
        if(r == me) continue;
        uint32_t * fn = c->let_scan_n.as<uint32_t>(), * fp = c->let_scan_p.as<uint32_t>();
        let_flags_kernel<<<let_grid(c, nn0), 256, 0, c->stream>>>(c->let_mask.as<unsigned>(), c->t_lk.as<int4>(), (int) nn0, r, fn, fp);
        SQ_TRY(sq_exclusive_scan_u32(c, fn, nn0, fn + nn0));
        SQ_TRY(sq_exclusive_scan_u32(c, fp, nn0, fp + nn0));
        double4 * pcm, * pcl; int4 * plk; double * ppos; float * pmass; int * proot;
        sect(c->let_send.as<unsigned char>() + soff[r], snn[r], snp[r], pcm, pcl, plk, ppos, pmass, proot);
        let_pack_kernel<<<let_grid(c, nn0), 256, 0, c->stream>>>(c->let_mask.as<unsigned>(), r, c->t_lk.as<int4>(), c->t_cm.as<double4>(), c->t_cl.as<double4>(),
                SP.as<double>(), SM.as<float>(), (int) nn0, fn, fp, pcm, pcl, plk, ppos, pmass);
        let_roots_kernel<<<let_grid(c, NL), 256, 0, c->stream>>>(c->d_topleaves.as<sq_topleaf>(), NL, me, c->d_treenode.as<int>(), c->let_mask.as<unsigned>(), r, fn, proot);
        c->kernel_launches += 3;
        LI.exported_nodes += snn[r]; LI.exported_particles += snp[r];
        LI.bytes_sent += (int64_t) peer_bytes(snn[r], snp[r]);
    }
    SQ_CUDA(cudaGetLastError());

    // ---- sizes of what every rank sends to every rank (all-gather of the rows)
    sq_buf & tab = c->let_roots;
    SQ_TRY(tab.reserve(sizeof(int64_t) * 2 * (size_t) R * (size_t) (R + 1)));
    std::vector<int64_t> row(2 * R, 0), all(2 * (size_t) R * R, 0);
    for(int r = 0; r < R; r++) { row[2 * r] = snn[r]; row[2 * r + 1] = snp[r]; }
    int64_t * d_row = tab.as<int64_t>(), * d_all = d_row + 2 * R;
    SQ_CUDA(cudaMemcpyAsync(d_row, row.data(), sizeof(int64_t) * 2 * R, cudaMemcpyHostToDevice, c->stream));
    SQ_TRY(sq_comm_allgather_i64(c, d_row, d_all, (size_t) 2 * R));
    SQ_CUDA(cudaMemcpyAsync(all.data(), d_all, sizeof(int64_t) * 2 * R * R, cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    std::vector<int64_t> rnn(R, 0), rnp(R, 0), roff(R, 0);
    size_t recv_total = 0;
    int64_t imp_n = 0, imp_p = 0;
    for(int s = 0; s < R; s++) {
        if(s == me) continue;
        rnn[s] = all[(size_t) s * 2 * R + 2 * me]; rnp[s] = all[(size_t) s * 2 * R + 2 * me + 1];
        roff[s] = (int64_t) recv_total; recv_total += peer_bytes(rnn[s], rnp[s]);
        imp_n += rnn[s]; imp_p += rnp[s];
    }
    if(nn0 + imp_n >= 0x5fffffff || n + imp_p >= ((int64_t) 1 << 31)) { sq_set_error("sq_tree_build_dist: local + imported tree exceeds the 32-bit index space"); return SQ_ERR_NODE_POOL; }
    SQ_TRY(c->let_recv.reserve(recv_total + 256));
    SQ_CUDA(cudaEventRecord(ev[14], c->stream));
    {
        std::vector<const unsigned char *> sptr(R, nullptr);
        std::vector<unsigned char *> rptr(R, nullptr);
        std::vector<size_t> sb(R, 0), rb(R, 0);
        for(int r = 0; r < R; r++) {
            if(r == me) continue;
            sptr[r] = c->let_send.as<unsigned char>() + soff[r]; sb[r] = peer_bytes(snn[r], snp[r]);
            rptr[r] = c->let_recv.as<unsigned char>() + roff[r]; rb[r] = peer_bytes(rnn[r], rnp[r]);
            LI.bytes_received += (int64_t) rb[r];
        }
        SQ_TRY(sq_comm_exchange(c, sptr, sb, rptr, rb));
    }
    SQ_CUDA(cudaEventRecord(ev[15], c->stream));

    // ---- receiver: append behind the local tree and particles, shift the links, hang the roots below the pseudo nodes
    const size_t ntotn = (size_t) (nn0 + imp_n + 1), ntotp = (size_t) (n + imp_p + 1);
    SQ_TRY(c->t_cm.grow_keep(sizeof(double4) * ntotn, sizeof(double4) * (size_t) nn0, c->stream));
    SQ_TRY(c->t_cl.grow_keep(sizeof(double4) * ntotn, sizeof(double4) * (size_t) nn0, c->stream));
    SQ_TRY(c->t_lk.grow_keep(sizeof(int4) * ntotn, sizeof(int4) * (size_t) nn0, c->stream));
    SQ_TRY(SP.grow_keep(sizeof(double) * 3 * ntotp, sizeof(double) * 3 * (size_t) n, c->stream));
    SQ_TRY(SM.grow_keep(sizeof(float) * ntotp, sizeof(float) * (size_t) n, c->stream));
    int64_t noff = nn0, poff = n;
    for(int s = 0; s < R; s++) {
        if(s == me) continue;
        double4 * pcm, * pcl; int4 * plk; double * ppos; float * pmass; int * proot;
        sect(c->let_recv.as<unsigned char>() + roff[s], rnn[s], rnp[s], pcm, pcl, plk, ppos, pmass, proot);
        if(rnn[s] > 0) {
            SQ_CUDA(cudaMemcpyAsync(c->t_cm.as<double4>() + noff, pcm, sizeof(double4) * (size_t) rnn[s], cudaMemcpyDeviceToDevice, c->stream));
            SQ_CUDA(cudaMemcpyAsync(c->t_cl.as<double4>() + noff, pcl, sizeof(double4) * (size_t) rnn[s], cudaMemcpyDeviceToDevice, c->stream));
            SQ_CUDA(cudaMemcpyAsync(c->t_lk.as<int4>() + noff, plk, sizeof(int4) * (size_t) rnn[s], cudaMemcpyDeviceToDevice, c->stream));
        }
        if(rnp[s] > 0) {
            SQ_CUDA(cudaMemcpyAsync(SP.as<double>() + 3 * poff, ppos, sizeof(double) * 3 * (size_t) rnp[s], cudaMemcpyDeviceToDevice, c->stream));
            SQ_CUDA(cudaMemcpyAsync(SM.as<float>() + poff, pmass, sizeof(float) * (size_t) rnp[s], cudaMemcpyDeviceToDevice, c->stream));
        }
        if(rnn[s] > 0) {
            let_relocate_kernel<<<let_grid(c, rnn[s]), 256, 0, c->stream>>>(c->t_lk.as<int4>(), noff, rnn[s], poff);
            let_attach_kernel<<<let_grid(c, NL), 256, 0, c->stream>>>(d_tasks.as<sq_task>(), s, proot, c->d_treenode.as<int>(), c->t_lk.as<int4>(), noff);
            c->kernel_launches += 2;
        }
        noff += rnn[s]; poff += rnp[s];
    }
    SQ_CUDA(cudaGetLastError());
    SQ_CUDA(cudaEventRecord(ev[7], c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    c->tree_numnodes = nn0 + imp_n;
    c->n_sources = n + imp_p;
    c->let_attached = true;
    LI.imported_nodes = imp_n; LI.imported_particles = imp_p;
    LI.build_ms = let_ms(ev[10], ev[11]); LI.moments_exchange_ms = let_ms(ev[11], ev[12]); LI.let_mark_ms = let_ms(ev[12], ev[14]);
    LI.let_exchange_ms = let_ms(ev[14], ev[15]); LI.let_attach_ms = let_ms(ev[15], ev[7]);
    if(info) info->numnodes = c->tree_numnodes;
    if(let) *let = LI;
    return SQ_OK;
}

extern "C" int sq_toptree_export_count(sq_ctx * c, int64_t * nexports, int64_t * nparticles_exporting)
{
    if(!c || !c->tree_built || !c->have_domain) { sq_set_error("sq_toptree_export_count: build a domain tree first"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaSetDevice(c->device));
    const sq_params & par = c->par;
    let_top_par P;
    P.box = par.BoxSize; P.half = 0.5 * par.BoxSize;
    P.rcut = par.TreeRcut * par.Asmth * (par.BoxSize / par.Nmesh); P.rcut2 = P.rcut * P.rcut;
    P.errtol = par.ErrTolForceAcc; P.relative = par.TreeUseBH == 0;
    P.bh2 = P.relative ? par.MaxBHOpeningAngle * par.MaxBHOpeningAngle : par.BHOpeningAngle * par.BHOpeningAngle;
    P.myrank = c->comm_rank;
    SQ_TRY(c->w_stats.reserve(64));
    SQ_CUDA(cudaMemsetAsync(c->w_stats.p, 0, 64, c->stream));
    if(c->n > 0) {
        let_toptree_count_kernel<<<let_grid(c, c->n), 256, 0, c->stream>>>(c->t_lk.as<int4>(), c->t_cm.as<double4>(), c->t_cl.as<double4>(), c->t_top.as<int>(),
                c->d_topnodes.as<sq_topnode>(), c->d_topleaves.as<sq_topleaf>(), c->pos.as<double>(), c->oldacc.as<double>(), (long long) c->n,
                c->w_stats.as<unsigned long long>(), P);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    unsigned long long h[2] = {0, 0};
    SQ_CUDA(cudaMemcpyAsync(h, c->w_stats.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    if(nexports) *nexports = (int64_t) h[0];
    if(nparticles_exporting) *nparticles_exporting = (int64_t) h[1];
    return SQ_OK;
}
