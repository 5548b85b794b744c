// The domain the multi-rank tree is built under: the reference's DomainDecomp (libgadget/domain.h:12-43).
//
//   sq_domain_set             installs topnode_data / topleaf_data / task_data arrays as the reference hands them over
//   sq_domain_from_histogram  pure host code: a top tree from a Peano-Hilbert key histogram, refined like
//                             domain_determine_global_toptree (domain.cpp:1040-1160: a node is split while it holds more
//                             than total / (DomainOverDecompositionFactor * NTask) of the cost, children in key order,
//                             StartKey / Shift as domain.cpp:849-868), leaves numbered in key order
//                             (domain_create_topleaves) and cut into NTask contiguous segments of equal cost
//                             (domain_assign_topleaves_balanced, domain.cpp:620-800, one segment per task)
//   sq_domain_histogram       key histogram of the uploaded particles on the device (+ ncclAllReduce over ranks)
//   sq_domain_owner           domain_get_topleaf (domain.h:68-75) per particle -> owning rank
// The full particle exchange (domain_exchange, exchange.hpp) stays with the host: SURVEY 8(f1).
#include "sq_common.cuh"
#include <algorithm>
#include <math.h>

int sq_comm_allreduce_sum_i64(sq_ctx * c, int64_t * dev, size_t n);
int sq_comm_allreduce_sum_f64(sq_ctx * c, double * dev, size_t n);

// ------------------------------------------------------------------------------------ host: top tree from a histogram
namespace {
struct hist_view {
    const int64_t * count; const double * cost; int level;
    // totals of the cells below the node (StartKey, Shift): cells [StartKey >> s, (StartKey + 2^Shift) >> s), s = 63 - 3 level
    void sum(uint64_t start, int shift, int64_t & n, double & w) const
    {
        const int s = 63 - 3 * level;
        const uint64_t a = start >> s, b = a + ((uint64_t) 1 << (shift - s));
        n = 0; w = 0;
        for(uint64_t i = a; i < b; i++) { n += count[i]; w += cost ? cost[i] : (double) count[i]; }
    }
};
}

extern "C" int sq_domain_from_histogram(const int64_t * count, const double * cost, int level, int nranks, int overdecomposition,
                                        sq_topnode * topnodes, sq_topleaf * topleaves, sq_task * tasks, int max_topnodes,
                                        int * ntopnodes, int * ntopleaves)
{
    if(!count || !topnodes || !topleaves || !tasks || !ntopnodes || !ntopleaves) { sq_set_error("sq_domain_from_histogram: null argument"); return SQ_ERR_ARG; }
    if(level < 0 || level > 7 || nranks < 1 || overdecomposition < 1 || max_topnodes < 1) { sq_set_error("sq_domain_from_histogram: level 0..7, nranks >= 1, overdecomposition >= 1"); return SQ_ERR_ARG; }
    hist_view H{count, cost, level};
    int64_t ntot; double wtot;
    H.sum(0, 63, ntot, wtot);
    const double limit = wtot / ((double) overdecomposition * nranks);
    std::vector<double> w(1, wtot);
    std::vector<int64_t> cnt(1, ntot);
    int nn = 1;
    topnodes[0].StartKey = 0; topnodes[0].Daughter = -1; topnodes[0].Shift = 63; topnodes[0].Leaf = -1; topnodes[0].pad_ = 0;
    // breadth first: a node is refined while it is too heavy, or while there are fewer leaves than ranks can share out
    for(int i = 0; i < nn; i++) {
        const int lvl = (63 - topnodes[i].Shift) / 3;
        const bool heavy = w[i] > limit && cnt[i] > 0;
        if(!heavy || lvl >= level) continue;
        if(nn + 8 > max_topnodes) { sq_set_error("sq_domain_from_histogram: more than %d top nodes needed", max_topnodes); return SQ_ERR_ARG; }
        topnodes[i].Daughter = nn;
        for(int j = 0; j < 8; j++) {
            sq_topnode & t = topnodes[nn + j];
            t.Shift = topnodes[i].Shift - 3;
            t.StartKey = topnodes[i].StartKey + ((uint64_t) j << t.Shift);
            t.Daughter = -1; t.Leaf = -1; t.pad_ = 0;
            int64_t n; double ww;
            H.sum(t.StartKey, t.Shift, n, ww);
            cnt.push_back(n); w.push_back(ww);
        }
        nn += 8;
    }
    // leaves in key order = depth first through the daughters (domain_create_topleaves)
    std::vector<int> order;
    {
        std::vector<int> stack(1, 0);
        while(!stack.empty()) {
            const int i = stack.back(); stack.pop_back();
            if(topnodes[i].Daughter < 0) order.push_back(i);
            else for(int j = 7; j >= 0; j--) stack.push_back(topnodes[i].Daughter + j);
        }
    }
    const int nl = (int) order.size();
    if(nl < nranks) { sq_set_error("Number of Topleaves is less than NTask (%d < %d): raise the histogram level", nl, nranks); return SQ_ERR_ARG; }
    // contiguous segments of equal cost, one per rank; every rank gets at least one leaf
    std::vector<double> pre(nl + 1, 0.0);
    for(int k = 0; k < nl; k++) pre[k + 1] = pre[k] + w[order[k]];
    int start = 0;
    for(int t = 0; t < nranks; t++) {
        int end;
        if(t == nranks - 1) end = nl;
        else {
            const double want = wtot * (double) (t + 1) / nranks;
            end = (int) (std::lower_bound(pre.begin(), pre.end(), want) - pre.begin());
            if(end > 0 && end <= nl && fabs(pre[end - 1] - want) < fabs(pre[end < nl ? end : nl] - want)) end--;
            if(end <= start) end = start + 1;
            if(end > nl - (nranks - 1 - t)) end = nl - (nranks - 1 - t);
        }
        tasks[t].StartLeaf = start; tasks[t].EndLeaf = end;
        for(int k = start; k < end; k++) {
            topleaves[k].Task = t; topleaves[k].topnode = order[k]; topleaves[k].treenode = -1;
            topnodes[order[k]].Leaf = k;
        }
        start = end;
    }
    *ntopnodes = nn; *ntopleaves = nl;
    return SQ_OK;
}

// ------------------------------------------------------------------------------------ host: the reference's leaf assignment
// domain_assign_topleaves_balanced (domain.cpp:620-778) restated: the top leaves, in Peano-Hilbert key order with their
// integer costs, are cut greedily into nranks * nsegment_per_task contiguous segments whose running total follows the
// mean (a leaf joins the open segment while less than half of it would overshoot; a segment is never empty; the last
// leaves are kept back so that every segment gets one), and consecutive segments fill one task after the other until
// the task holds its share; when every task has been served and leaves remain, the means are recomputed from what is
// left and the tasks are served again.  task[k] = task of leaf k.  Pinned against the reference's own function for
// random costs and task counts (oracle/ref_domain.cpp, tests/test_domain_cpu.py): equal for every leaf.
extern "C" int sq_domain_assign_balanced(const int64_t * cost, int ntopleaves, int nranks, int nsegment_per_task, int32_t * task)
{
    if(!cost || !task || ntopleaves < 1 || nranks < 1 || nsegment_per_task < 1) { sq_set_error("sq_domain_assign_balanced: bad argument"); return SQ_ERR_ARG; }
    const int nseg = nranks * nsegment_per_task;
    if(ntopleaves < nseg) { sq_set_error("sq_domain_assign_balanced: %d top leaves cannot fill %d segments", ntopleaves, nseg); return SQ_ERR_ARG; }
    int64_t total = 0;
    for(int k = 0; k < ntopleaves; k++) {
        if(cost[k] < 0) { sq_set_error("sq_domain_assign_balanced: negative cost"); return SQ_ERR_ARG; }
        total += cost[k];
    }
    int64_t unassigned = total;              // cost of the leaves behind the closed segments
    double seg_share = 1.0 * total / nseg;   // what a segment / a task should hold in this pass over the tasks
    double task_share = 1.0 * total / nranks;
    int leaf = 0, seg = 0, t = 0, passes = 0;
    int64_t open = 0, task_load = 0;         // cost in the open segment / in the task being filled
    while(passes < ntopleaves) {
        bool take, close;
        if(leaf == ntopleaves) { take = false; close = true; }
        else if(ntopleaves - leaf == nseg - seg) { take = true; close = true; }     // one leaf per segment still to come
        else {
            const int64_t done = (total - unassigned) + open;
            take = (seg_share * (seg + 1) - done > 0.5 * cost[leaf]) || open == 0;
            close = !take;
        }
        if(take) {
            open += cost[leaf];
            task[leaf++] = t;
        }
        if(!close) continue;
        task_load += open;
        if((task_share - task_load < 0.5 * seg_share) || (nseg - seg <= nranks - t)) {
            task_load = 0;
            t++;
        }
        unassigned -= open;
        open = 0;
        seg++;
        if(t == nranks) {
            t = 0;
            seg_share = 1.0 * unassigned / nseg;
            task_share = 1.0 * unassigned / nranks;
            passes++;
        }
        if(leaf == ntopleaves) break;
    }
    if(seg < nseg || unassigned != 0) { sq_set_error("sq_domain_assign_balanced: %d of %d segments, %lld cost left", seg, nseg, (long long) unassigned); return SQ_ERR_ARG; }
    return SQ_OK;
}

// ------------------------------------------------------------------------------------ install / query
static int sq_domain_check(const sq_topnode * tn, int ntn, const sq_topleaf * tl, int ntl, const sq_task * tk, int ntk, int * maxlevel)
{
    if(ntn < 1 || ntl < 1 || ntk < 1) return 1;
    if(tn[0].Shift != 63 || tn[0].StartKey != 0) return 2;
    int ml = 0, leaves = 0;
    for(int i = 0; i < ntn; i++) {
        if(tn[i].Shift < 0 || tn[i].Shift > 63 || (63 - tn[i].Shift) % 3) return 3;
        const int lvl = (63 - tn[i].Shift) / 3;
        if(lvl > ml) ml = lvl;
        if(tn[i].Daughter >= 0) {
            if(tn[i].Daughter + 8 > ntn || tn[i].Shift < 3) return 4;
            for(int j = 0; j < 8; j++) {
                const sq_topnode & d = tn[tn[i].Daughter + j];
                if(d.Shift != tn[i].Shift - 3 || d.StartKey != tn[i].StartKey + ((uint64_t) j << d.Shift)) return 5;
            }
        } else {
            if(tn[i].Leaf < 0 || tn[i].Leaf >= ntl) return 6;
            leaves++;
        }
    }
    if(leaves != ntl) return 7;
    if(ml > SQ_MAXLEVEL - 1) return 8;
    for(int k = 0; k < ntl; k++)
        if(tl[k].Task < 0 || tl[k].Task >= ntk) return 9;
    for(int t = 0; t < ntk; t++) {
        if(tk[t].StartLeaf < 0 || tk[t].EndLeaf > ntl || tk[t].EndLeaf < tk[t].StartLeaf) return 10;
        for(int k = tk[t].StartLeaf; k < tk[t].EndLeaf; k++)
            if(tl[k].Task != t) return 11;
    }
    *maxlevel = ml;
    return 0;
}

// the consistency checks sq_domain_set applies, without a context (pure host code): 0 = accepted, else the failed check
extern "C" int sq_domain_validate(const sq_topnode * topnodes, int ntopnodes, const sq_topleaf * topleaves, int ntopleaves,
                                  const sq_task * tasks, int ntasks, int * maxlevel)
{
    if(!topnodes || !topleaves || !tasks) { sq_set_error("sq_domain_validate: null argument"); return SQ_ERR_ARG; }
    int ml = 0;
    const int bad = sq_domain_check(topnodes, ntopnodes, topleaves, ntopleaves, tasks, ntasks, &ml);
    if(bad) { sq_set_error("inconsistent DomainDecomp (check %d)", bad); return SQ_ERR_ARG; }
    if(maxlevel) *maxlevel = ml;
    return SQ_OK;
}

extern "C" int sq_domain_set(sq_ctx * c, const sq_topnode * topnodes, int ntopnodes, const sq_topleaf * topleaves, int ntopleaves,
                             const sq_task * tasks, int ntasks)
{
    if(!c) return SQ_ERR_ARG;
    SQ_CUDA(cudaSetDevice(c->device));
    c->tree_built = false;
    if(!topnodes || ntopnodes == 0) {
        c->have_domain = false;
        c->dom_nodes.clear(); c->dom_leaves.clear(); c->dom_tasks.clear();
        return SQ_OK;
    }
    if(!topleaves || !tasks) { sq_set_error("sq_domain_set: null argument"); return SQ_ERR_ARG; }
    if(ntasks != c->nranks) { sq_set_error("sq_domain_set: the domain is for %d tasks, this context is rank %d of %d", ntasks, c->comm_rank, c->nranks); return SQ_ERR_ARG; }
    int ml = 0;
    const int bad = sq_domain_check(topnodes, ntopnodes, topleaves, ntopleaves, tasks, ntasks, &ml);
    if(bad) { sq_set_error("sq_domain_set: inconsistent DomainDecomp (check %d)", bad); return SQ_ERR_ARG; }
    c->dom_nodes.assign(topnodes, topnodes + ntopnodes);
    c->dom_leaves.assign(topleaves, topleaves + ntopleaves);
    c->dom_tasks.assign(tasks, tasks + ntasks);
    c->dom_maxlevel = ml;
    SQ_TRY(c->d_topnodes.reserve(sizeof(sq_topnode) * (size_t) ntopnodes));
    SQ_TRY(c->d_topleaves.reserve(sizeof(sq_topleaf) * (size_t) ntopleaves));
    SQ_TRY(c->d_treenode.reserve(sizeof(int) * (size_t) ntopleaves));
    SQ_CUDA(cudaMemcpyAsync(c->d_topnodes.p, c->dom_nodes.data(), sizeof(sq_topnode) * (size_t) ntopnodes, cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->d_topleaves.p, c->dom_leaves.data(), sizeof(sq_topleaf) * (size_t) ntopleaves, cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaMemsetAsync(c->d_treenode.p, 0xff, sizeof(int) * (size_t) ntopleaves, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    c->have_domain = true;
    return SQ_OK;
}

extern "C" int sq_domain_get_treenodes(sq_ctx * c, int32_t * treenode)
{
    if(!c || !treenode || !c->have_domain || !c->tree_built) { sq_set_error("sq_domain_get_treenodes: no domain tree"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaMemcpyAsync(treenode, c->d_treenode.p, sizeof(int) * c->dom_leaves.size(), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    return SQ_OK;
}

// ------------------------------------------------------------------------------------ device: histogram, owners
__global__ void sq_dom_hist_kernel(const uint64_t * __restrict__ keys, const float * __restrict__ wts, long long n, int shift,
                                   unsigned long long * __restrict__ count, double * __restrict__ cost)
{
    for(long long i = blockIdx.x * (long long) blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const unsigned long long cell = keys[i] >> shift;
        atomicAdd(&count[cell], 1ull);
        if(cost) atomicAdd(&cost[cell], wts ? (double) wts[i] : 1.0);
    }
}

// domain_get_topleaf, domain.h:68-75
__device__ __forceinline__ int sq_dom_topleaf(const sq_topnode * __restrict__ tn, uint64_t key)
{
    int no = 0;
    while(tn[no].Daughter >= 0)
        no = tn[no].Daughter + (int) ((key - tn[no].StartKey) >> (tn[no].Shift - 3));
    return tn[no].Leaf;
}

__global__ void sq_dom_owner_kernel(const uint64_t * __restrict__ keys, long long n, const sq_topnode * __restrict__ tn,
                                    const sq_topleaf * __restrict__ tl, int32_t * __restrict__ owner)
{
    for(long long i = blockIdx.x * (long long) blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x)
        owner[i] = tl[sq_dom_topleaf(tn, keys[i])].Task;
}

static int dom_grid(const sq_ctx * c, long long n)
{
    long long need = (n + 255) / 256, cap = (long long) c->num_sms * 8;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}

extern "C" int sq_domain_histogram(sq_ctx * c, int level, int64_t * count, double * cost, const float * weights)
{
    if(!c || !count || level < 0 || level > 7) { sq_set_error("sq_domain_histogram: bad argument"); return SQ_ERR_ARG; }
    if(!c->have_params) { sq_set_error("sq_domain_histogram: call sq_set_params first"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaSetDevice(c->device));
    SQ_TRY(sq_keys_run(c));
    const size_t ncell = (size_t) 1 << (3 * level);
    sq_buf hc, hw, dw;
    int rc = SQ_OK;
    do {
        if((rc = hc.reserve(sizeof(int64_t) * ncell))) break;
        if(cost && (rc = hw.reserve(sizeof(double) * ncell))) break;
        cudaError_t e = cudaMemsetAsync(hc.p, 0, sizeof(int64_t) * ncell, c->stream);
        if(e == cudaSuccess && cost) e = cudaMemsetAsync(hw.p, 0, sizeof(double) * ncell, c->stream);
        const float * w = nullptr;
        if(e == cudaSuccess && cost && weights && c->n > 0) {
            if((rc = dw.reserve(sizeof(float) * (size_t) c->n))) break;
            e = cudaMemcpyAsync(dw.p, weights, sizeof(float) * (size_t) c->n, cudaMemcpyDefault, c->stream);
            w = dw.as<float>();
        }
        if(e != cudaSuccess) { sq_set_error("sq_domain_histogram: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; break; }
        if(c->n > 0) {
            sq_dom_hist_kernel<<<dom_grid(c, c->n), 256, 0, c->stream>>>(c->keys_in.as<uint64_t>(), w, (long long) c->n, 63 - 3 * level,
                    hc.as<unsigned long long>(), cost ? hw.as<double>() : nullptr);
            c->kernel_launches++;
        }
        if(c->nranks > 1) {
            if((rc = sq_comm_allreduce_sum_i64(c, hc.as<int64_t>(), ncell))) break;
            if(cost && (rc = sq_comm_allreduce_sum_f64(c, hw.as<double>(), ncell))) break;
        }
        e = cudaMemcpyAsync(count, hc.p, sizeof(int64_t) * ncell, cudaMemcpyDeviceToHost, c->stream);
        if(e == cudaSuccess && cost) e = cudaMemcpyAsync(cost, hw.p, sizeof(double) * ncell, cudaMemcpyDeviceToHost, c->stream);
        if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if(e != cudaSuccess) { sq_set_error("sq_domain_histogram: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; break; }
    } while(0);
    hc.release(); hw.release(); dw.release();
    return rc;
}

extern "C" int sq_domain_owner(sq_ctx * c, int32_t * owner)
{
    if(!c || !owner || !c->have_domain) { sq_set_error("sq_domain_owner: no domain installed"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaSetDevice(c->device));
    SQ_TRY(sq_keys_run(c));
    sq_buf o;
    SQ_TRY(o.reserve(sizeof(int32_t) * (size_t) (c->n + 1)));
    int rc = SQ_OK;
    if(c->n > 0) {
        sq_dom_owner_kernel<<<dom_grid(c, c->n), 256, 0, c->stream>>>(c->keys_in.as<uint64_t>(), (long long) c->n, c->d_topnodes.as<sq_topnode>(),
                c->d_topleaves.as<sq_topleaf>(), o.as<int32_t>());
        c->kernel_launches++;
        cudaError_t e = cudaMemcpyAsync(owner, o.p, sizeof(int32_t) * (size_t) c->n, cudaMemcpyDeviceToHost, c->stream);
        if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if(e != cudaSuccess) { sq_set_error("sq_domain_owner: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; }
    }
    o.release();
    return rc;
}
