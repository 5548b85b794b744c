// shenqi_b200 internal header: context, device buffers, error plumbing, warp helpers.
// sm_100a only.  No Thrust/CUB, no CPU fallback.
#pragma once
#include <cuda_runtime.h>
#include <cufft.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/shenqi_b200.h"

#define SQ_WARP 32
#define SQ_FULL 0xffffffffu
#define SQ_NMAXCHILD 8            /* libgadget/forcetree.h:13 */
#define SQ_NODEFULL (1 << 16)     /* libgadget/forcetree.h:14 */
#define SQ_MAXLEVEL 21            /* BITS_PER_DIMENSION, libgadget/utils/peano.h:9 */
#define SQ_NTAB 512               /* NGRAVTAB, libgadget/gravity.h:35 */

void sq_set_error(const char * fmt, ...);

#define SQ_CUDA(call)                                                                         \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if(e_ != cudaSuccess) {                                                               \
            sq_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return SQ_ERR_CUDA;                                                               \
        }                                                                                     \
    } while(0)

#define SQ_CUFFT(call)                                                                        \
    do {                                                                                      \
        cufftResult r_ = (call);                                                              \
        if(r_ != CUFFT_SUCCESS) {                                                             \
            sq_set_error("%s:%d: %s -> cufft error %d", __FILE__, __LINE__, #call, (int) r_); \
            return SQ_ERR_CUDA;                                                               \
        }                                                                                     \
    } while(0)

#define SQ_TRY(call) do { int rc_ = (call); if(rc_ != SQ_OK) return rc_; } while(0)

// Growable device buffer owned by the context (the library owns its device pools).
struct sq_buf {
    void * p = nullptr;
    size_t cap = 0;
    // grow to `bytes`, keeping the first `keep` bytes (device-to-device copy on `stream`)
    int grow_keep(size_t bytes, size_t keep, cudaStream_t stream)
    {
        if(bytes <= cap) return SQ_OK;
        void * q = nullptr;
        size_t want = bytes + bytes / 16 + 256;
        SQ_CUDA(cudaMalloc(&q, want));
        if(p && keep > 0) SQ_CUDA(cudaMemcpyAsync(q, p, keep < cap ? keep : cap, cudaMemcpyDeviceToDevice, stream));
        if(p) { SQ_CUDA(cudaStreamSynchronize(stream)); cudaFree(p); }
        p = q; cap = want;
        return SQ_OK;
    }
    int reserve(size_t bytes)
    {
        if(bytes <= cap) return SQ_OK;
        if(p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 16 + 256;
        SQ_CUDA(cudaMalloc(&p, want));
        cap = want;
        return SQ_OK;
    }
    void release() { if(p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T * as() const { return (T *) p; }
};

// Packed device node, three 128-bit-aligned streams (DESIGN.md "data layout"):
//   cm[i]  = {cofm.x, cofm.y, cofm.z, mass}      32 B   (two 128-bit loads)
//   cl[i]  = {center.x, center.y, center.z, len}  32 B
//   lk[i]  = {first, count, sibling, info}        16 B   (one 128-bit load)
//     internal node: first = index of first child, count = number of children (contiguous, Morton order)
//     leaf:          first = first particle (tree order), count = particles (<= 8)
//     info = level | type << 8 | toplevel << 12 | internaltop << 13 | dependslocal << 14 | octant << 16 | child octant mask << 24
struct sq_tree_dev {
    double4 * cm = nullptr;
    double4 * cl = nullptr;
    int4 * lk = nullptr;
    int * parent = nullptr;
    unsigned char * rot = nullptr;   // Peano-Hilbert orientation of the node's cell
};
#define SQ_INFO_LEVEL(i) ((i) & 0xff)
#define SQ_INFO_TYPE(i) (((i) >> 8) & 3)      /* 0 particle node, 1 node node, 2 pseudo */
#define SQ_INFO_TOP(i) (((i) >> 12) & 1)
#define SQ_INFO_ITOP(i) (((i) >> 13) & 1)
#define SQ_INFO_LOCAL(i) (((i) >> 14) & 1)
#define SQ_INFO_OCT(i) (((i) >> 16) & 7)
#define SQ_INFO_CHILDMASK(i) (((unsigned) (i)) >> 24)   /* internal nodes: octants (x + 2y + 4z) that have a child; children are stored in this order */

struct sq_ctx {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[16] = {};
    cudaStream_t copy_stream = nullptr;   // sq_gravity_step: host copies overlapped with the kernels
    cudaEvent_t evx[4] = {};
    sq_params par = {};
    bool have_params = false;
    sq_walk_raw raw = {};
    bool have_raw = false;
    float raw_ft[SQ_NTAB], raw_pt[SQ_NTAB];
    uint64_t raw_serial = 0, w_table_sig = 0;     // device split table: rebuilt only when its inputs change
    int64_t kernel_launches = 0;
    sq_timings tm = {};

    // ---- particles, upload order
    int64_t n = 0;
    sq_buf pos_in, mass_in, id_in, type_in, oldacc_in, keys_in, oldacc_tmp, stage_a, stage_b;
    bool have_types = false;
    // IsGarbage || (Swallowed && Type == 5) per uploaded particle (forcetree.cpp:805, slotsmanager.cpp:428-432): such
    // particles get the largest key, sort behind every valid one and stay out of the tree, the walk and the mesh
    sq_buf skip_in;
    bool have_skip = false;
    int64_t n_valid = 0;             // particles in front of the garbage after the sort
    bool walk_all = false;           // the last walk had no explicit target list
    // ---- particles, PH-sorted order
    sq_buf pos, mass, oldacc, keys, perm, rank;   // perm[sorted] = upload idx; rank[upload] = sorted pos
    sq_buf sort_tmp_keys, sort_tmp_vals, sort_hist, scan_tmp, flag_tmp;
    bool sorted = false;
    int geo_mismatch = 0;

    // ---- tree
    sq_buf t_cm, t_cl, t_lk, t_parent, t_rot, t_lvl, t_counters;
    sq_buf tree_keys, tree_order, tree_pos, tree_mass;   // only when the tree order differs from particle order
    sq_buf active_sorted;            // sorted positions of the active set the tree was built from
    sq_buf act_tmp;                  // the caller's active list on the device
    int64_t tree_np = 0;             // particles in the tree
    int64_t tree_maxnodes = 0, tree_numnodes = 0;
    int tree_maxlevel = 0;
    double tree_total_mass = 0;      // root monopole mass (mass unit of the walk's FP32 tiers)
    int lvl_start[SQ_MAXLEVEL + 3] = {};
    bool tree_built = false, tree_full = false, tree_identity = true;

    // ---- walk
    sq_buf w_accel, w_pot, w_stats, w_targets, w_table, w_out_accel, w_out_pot, w_counter, w_icm, w_icl, w_ipart;
    sq_buf w_gcost;                  // interactions per target group of the last walk
    int walk_group = 32;
    bool walk_done = false;
    int64_t walk_ntargets = 0;

    // ---- PM
    sq_buf pm_real[4], pm_cplx[4], pm_out;
    cufftHandle plan_r2c = 0, plan_c2r = 0;
    int plan_nmesh = 0;
    bool pm_done = false;

    // ---- multi-GPU (sq_dist.cu)
    bool own_stream = true;
    int64_t range_first = 0, range_count = -1;      // sq_tree_walk_range: targets = this range of the sorted set
    cufftHandle slab_r2c = 0, slab_c2r = 0, slab_x = 0;
    int slab_N = 0, slab_P = 0;
    sq_buf slab_tab, route_tmp;
    sq_buf step_tmp;                 // sq_step.cu: error counter, compaction block counts

    // ---- multi-rank tree (sq_domain.cu, sq_let.cu): the reference's DomainDecomp + raw NCCL
    void * nccl_comm = nullptr;      // ncclComm_t
    void * loop_hub = nullptr;       // test transport: ranks as host threads of one process (sq_let.cu)
    int comm_rank = 0, nranks = 1;
    std::vector<sq_topnode> dom_nodes;
    std::vector<sq_topleaf> dom_leaves;
    std::vector<sq_task> dom_tasks;
    bool have_domain = false;
    int dom_maxlevel = 0;            // deepest top-node level
    sq_buf d_leafrange;              // int2 per top leaf: (first particle, count) of a local top leaf in the sorted order
    sq_buf d_topnodes, d_topleaves, d_treenode, t_top;   // device copies; treenode[leaf] = node of that top leaf; t_top[node] = top node index or -1
    sq_buf let_leafdata, let_mask, let_scan_n, let_scan_p, let_roots, let_send, let_recv, let_boxes;
    int64_t n_local_nodes = 0;       // nodes of this rank's own build (imports are appended behind them)
    int64_t n_sources = 0;           // particles the walk can read as sources: n local + imported
    bool let_attached = false;
    double walk_munit = 0;           // mass unit override for the walk's FP32 tiers (0: total / count of the local tree)
};

static inline int sq_div_up(int64_t a, int64_t b) { return (int) ((a + b - 1) / b); }

#ifdef __CUDACC__
__device__ __forceinline__ unsigned sq_lane() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned sq_lanemask_lt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

__device__ __forceinline__ int sq_warp_excl_scan(int v, int & total)
{
    int x = v;
    #pragma unroll
    for(int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(SQ_FULL, x, o);
        if((int) sq_lane() >= o) x += y;
    }
    total = __shfl_sync(SQ_FULL, x, 31);
    return x - v;
}
// Lanes of `valid` (the lanes that call) holding the same value of an up-to-8-bit key.  Eight ballots: MATCH.ANY is slow on
// sm_100 (A/B on B200: radix sort 3.15 -> 2.04 ms, CIC paint 2.19 -> 1.19 ms after it was replaced).
__device__ __forceinline__ unsigned sq_match8(unsigned valid, unsigned key8)
{
    unsigned peers = valid;
    #pragma unroll
    for(int b = 0; b < 8; b++) {
        const bool bit = (key8 >> b) & 1u;
        const unsigned v = __ballot_sync(valid, bit);
        peers &= bit ? v : ~v;
    }
    return peers;
}
// Lanes of the warp (all 32 take part) whose 64-bit value equals this lane's: grouped by an 8-bit hash first (sq_match8),
// then every hash group is split by comparing with its first unassigned lane until every lane has found its group
// (one round unless two different values share a hash).
__device__ __forceinline__ unsigned sq_match64(long long val, unsigned hash8)
{
    const unsigned same_hash = sq_match8(SQ_FULL, hash8);
    unsigned todo = SQ_FULL, mine = 0;
    const unsigned me = 1u << sq_lane();
    while(todo) {
        const unsigned cand = same_hash & todo;
        const int leader = cand ? __ffs(cand) - 1 : 0;
        const long long lead = __shfl_sync(SQ_FULL, val, leader);
        const bool eq = (todo & me) && lead == val;
        const unsigned done = __ballot_sync(SQ_FULL, eq);
        if(eq) mine = done & same_hash;
        todo &= ~done;
    }
    return mine;
}
// libgadget/partmanager.h:99, evaluated exactly as the macro does
__device__ __forceinline__ double sq_nearest(double x, double box, double half)
{
    return (x > half) ? (x - box) : ((x < -half) ? (x + box) : x);
}
#endif

// stage entry points (implemented in the .cu files)
int sq_keys_run(sq_ctx * c);
int sq_geokeys_run(sq_ctx * c, const double * pos, int64_t n, uint64_t * gkeys);
int sq_sort_pairs(sq_ctx * c, uint64_t * keys, uint32_t * vals, int64_t n, int begin_bit, int end_bit);
int sq_exclusive_scan_u32(sq_ctx * c, uint32_t * data, int64_t n, uint32_t * total_out_dev);
int sq_gather_run(sq_ctx * c);
int sq_active_to_sorted(sq_ctx * c, const int32_t * active, int64_t nactive, sq_buf * out, int64_t * count);
int sq_tree_run(sq_ctx * c, const int32_t * active, int64_t nactive, sq_tree_info * info);
int sq_tree_moments_top(sq_ctx * c, int ls, int le);
void sq_comm_destroy(sq_ctx * c);
int sq_tree_export(sq_ctx * c, void * host_nodes, int64_t firstnode, int upload_order, int64_t lastnode);
int sq_walk_run(sq_ctx * c, const int32_t * active, int64_t nactive, sq_walk_stats * stats);
int sq_pm_run(sq_ctx * c, bool paint_only);
void sq_pm_destroy(sq_ctx * c);
int sq_pm_gradient(sq_ctx * c, const double * phi, int nout, int xwrap, double * fx, double * fy, double * fz);
void sq_dist_destroy(sq_ctx * c);
int sq_walk_range_targets(sq_ctx * c, sq_buf * out);
