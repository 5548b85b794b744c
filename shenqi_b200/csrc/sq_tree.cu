// Force tree: oct-tree over Peano-Hilbert-sorted particles + monopole moments.
//
// Replaces force_tree_create_nodes / add_particle_to_tree / create_new_node_layer /
// merge_partial_force_trees (libgadget/forcetree.cpp:726-859, 480-519, 369-476, 525-649) and
// force_update_node_parallel/_recursive (forcetree.cpp:1016-1142).  Not a translation: the reference
// inserts particles one by one into a linked tree; here particles are already in key order, so every
// cell is a contiguous key range and the tree is grown level by level, one 8-lane team per cell:
//   - a cell with <= 8 particles is a leaf (NMAXCHILD, forcetree.h:13) and gets its moments at once;
//   - otherwise the 8 lanes binary-search the 8 digit boundaries of the next key digit, ballot the
//     non-empty octants, and the warp claims the child slots with ONE atomicAdd (warp scan of the
//     teams' child counts).  Empty children are never created (the reference prunes them afterwards,
//     forcetree.cpp:1031-1048).  Children are laid out contiguously in the reference's octant order
//     x + 2y + 4z (forcetree.cpp:277-283), so a parent's children are one coalesced 128-bit-load run.
// The resulting topology is the one the reference's insertion produces: a cell is split iff it holds
// more than 8 particles, independent of insertion order.  Centres follow the reference's FP64
// recurrence (centre +- len/4) so the NODE image is bit-identical.
// Internal moments are then reduced bottom-up, level by level, 8 lanes per node with shuffles.
#include "sq_common.cuh"

const unsigned char * sq_phtab_device();
int sq_iota(sq_ctx * c, uint32_t * v, int64_t n);

// counters: [0] flags (bit0 geo mismatch, bit1 bad mass), [1] next free node, [2] error flags (bit0 pool overflow, bit1 depth)
#define CNT_FLAGS 0
#define CNT_NEXT 1
#define CNT_ERR 2

__global__ void sq_tree_root_kernel(double4 * cm, double4 * cl, int4 * lk, int * parent, unsigned char * rot, int * lvl, int * cnt,
                                    int np, double box, int * topidx)
{
    if(topidx) topidx[0] = 0;      // the root is top node 0 of the domain (force_tree_create_topnodes, forcetree.cpp:654-686)
    // forcetree.cpp:653-686: root cube 1.001*Box centred on Box/2; the trivial domain's single top leaf
    cl[0] = make_double4(box / 2., box / 2., box / 2., box * 1.001);
    cm[0] = make_double4(0, 0, 0, 0);
    lk[0] = make_int4(0, np, -1, 0 | (1 << 12) | (1 << 14));   // level 0, TopLevel, DependsOnLocalMass
    parent[0] = -1;
    rot[0] = 0;
    lvl[0] = 0;
    lvl[1] = 1;
    cnt[CNT_NEXT] = 1;
    cnt[CNT_ERR] = 0;
}

__global__ void sq_tree_advance_kernel(int * lvl, const int * cnt, int L, int maxnodes)
{
    int v = cnt[CNT_NEXT];
    lvl[L + 2] = v < maxnodes ? v : maxnodes;
}

// One level of the top-down build.  8 lanes per node.
__global__ void __launch_bounds__(256) sq_tree_split_kernel(const uint64_t * __restrict__ keys, const double * __restrict__ pos,
                                                            const float * __restrict__ mass, double4 * __restrict__ cm,
                                                            double4 * __restrict__ cl, int4 * __restrict__ lk, int * __restrict__ parent,
                                                            unsigned char * __restrict__ rot, const int * __restrict__ lvl, int * __restrict__ cnt,
                                                            int L, int maxnodes, const unsigned char * __restrict__ phtab,
                                                            const sq_topnode * __restrict__ tn, const sq_topleaf * __restrict__ tl,
                                                            int * __restrict__ treenode, int * __restrict__ topidx, int myrank, int2 * __restrict__ leafrange)
{
    // With a domain (tn != NULL) the top of the tree is the domain's top tree, force_create_node_for_topnode
    // (forcetree.cpp:868-933): a node that is an internal top node gets all 8 children whatever it holds (TopLevel,
    // never pruned, forcetree.cpp:1033-1041), the child whose key digit is `sub` being top node Daughter + sub; a top
    // leaf of another rank is a pseudo node (no particles here); a top leaf of this rank is built down as usual and
    // its node number recorded (TopLeaves[].treenode).
    __shared__ unsigned char rot1[384], sub1[384];
    for(int i = threadIdx.x; i < 384; i += blockDim.x) {
        rot1[i] = phtab[6144 + i];
        sub1[i] = phtab[6528 + i];
    }
    __syncthreads();
    if(cnt[CNT_ERR] & 1) return;     // pool overflowed on an earlier level: slots past it were never written
    const int ls = lvl[L], le = lvl[L + 1];
    const unsigned lane = sq_lane();
    const int sg = lane >> 3, sl = lane & 7;
    const int teams = (gridDim.x * blockDim.x) >> 3;
    const int team0 = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int shift = L < SQ_MAXLEVEL ? 3 * (SQ_MAXLEVEL - 1 - L) : 0;
    // all 4 teams of a warp iterate together so the warp-level collectives stay converged
    for(int base_node = ls + (team0 - sg); base_node < le; base_node += teams) {
        const int i = base_node + sg;
        const bool live = i < le;
        int4 link = make_int4(0, 0, -1, 0);
        if(live) link = lk[i];
        const int first = link.x, count = link.y;
        bool forced = false, pseudo = false;
        int daughter = -1, wflags = link.w;
        if(tn && live) {
            const int ti = topidx[i];
            if(ti >= 0) {
                const sq_topnode T = tn[ti];
                if(T.Daughter >= 0) { forced = true; daughter = T.Daughter; }
                else {
                    pseudo = tl[T.Leaf].Task != myrank;
                    wflags |= pseudo ? (2 << 8) : (1 << 14);               // PSEUDO_NODE_TYPE / DependsOnLocalMass
                    if(sl == 0) {
                        treenode[T.Leaf] = i;
                        leafrange[T.Leaf] = make_int2(first, pseudo ? 0 : count);
                        if(pseudo && count > 0) atomicOr(&cnt[CNT_ERR], 4);   // a particle of another rank's top leaf
                        lk[i] = make_int4(first, pseudo ? 0 : count, link.z, wflags);
                    }
                }
            }
        }
        const bool split = live && !pseudo && (forced || count > SQ_NMAXCHILD);
        int nchild = 0, slot = 0, s_oct = 0, pix = 0, start = 0, ccount = 0;
        unsigned rotn = 0;
        if(live && !split) {
            // leaf: moments now (forcetree.cpp:946-952, 984-1003)
            double m = 0, mx = 0, my = 0, mz = 0;
            if(sl < count && !pseudo) {
                const int p = first + sl;
                m = (double) mass[p];
                mx = m * pos[3 * (int64_t) p];
                my = m * pos[3 * (int64_t) p + 1];
                mz = m * pos[3 * (int64_t) p + 2];
            }
            const unsigned tm = 0xffu << (sg * 8);
            #pragma unroll
            for(int o = 4; o >= 1; o >>= 1) {
                m += __shfl_xor_sync(tm, m, o);
                mx += __shfl_xor_sync(tm, mx, o);
                my += __shfl_xor_sync(tm, my, o);
                mz += __shfl_xor_sync(tm, mz, o);
            }
            if(sl == 0) {
                const double4 c4 = cl[i];
                if(m > 0) cm[i] = make_double4(mx / m, my / m, mz / m, m);
                else cm[i] = make_double4(c4.x, c4.y, c4.z, 0.0);
            }
        }
        if(split) {
            if(L >= SQ_MAXLEVEL) {
                if(sl == 0) atomicOr(&cnt[CNT_ERR], 2);
            } else {
                rotn = rot[i];
                // lower bound of digit >= sl inside [first, first+count)
                int lo = first, hi = first + count;
                if(sl == 0) hi = first;
                while(lo < hi) {
                    const int mid = lo + ((hi - lo) >> 1);
                    const unsigned d = (unsigned) (keys[mid] >> shift) & 7u;
                    if(d < (unsigned) sl) lo = mid + 1; else hi = mid;
                }
                start = lo;
            }
        }
        // next team-lane's start (lane 7: end of the range)
        {
            int nxt = __shfl_down_sync(SQ_FULL, start, 1);
            if(sl == 7) nxt = first + count;
            ccount = split && L < SQ_MAXLEVEL ? nxt - start : 0;
        }
        const bool exists = ccount > 0 || (forced && split && L < SQ_MAXLEVEL);
        if(exists) {
            // octant whose key digit is sl in this orientation; reference child index = x + 2y + 4z
            #pragma unroll
            for(int p = 0; p < 8; p++)
                if(sub1[rotn * 8 + p] == sl) pix = p;
            s_oct = ((pix >> 2) & 1) | (((pix >> 1) & 1) << 1) | ((pix & 1) << 2);
        }
        unsigned mm = exists ? (1u << s_oct) : 0u;
        #pragma unroll
        for(int o = 4; o >= 1; o >>= 1)
            mm |= __shfl_xor_sync(SQ_FULL, mm, o);
        nchild = __popc(mm);
        slot = __popc(mm & ((1u << s_oct) - 1u));
        // one atomic per warp: exclusive scan of the 4 teams' child counts
        int want = (sl == 0) ? nchild : 0;
        int wtot;
        int off = sq_warp_excl_scan(want, wtot);
        int wbase = 0;
        if(wtot > 0) {
            if(lane == 0) wbase = atomicAdd(&cnt[CNT_NEXT], wtot);
            wbase = __shfl_sync(SQ_FULL, wbase, 0);
        }
        const int tbase = wbase + __shfl_sync(SQ_FULL, off, sg * 8);
        if(nchild > 0) {
            if(tbase + nchild > maxnodes) {
                if(sl == 0) atomicOr(&cnt[CNT_ERR], 1);
            } else {
                if(exists) {
                    const int c = tbase + slot;
                    const double4 pc = cl[i];
                    const double q = 0.25 * pc.w;
                    const int sib = (slot + 1 < nchild) ? (c + 1) : link.z;
                    cl[c] = make_double4((s_oct & 1) ? __dadd_rn(pc.x, q) : __dadd_rn(pc.x, -q),
                                         (s_oct & 2) ? __dadd_rn(pc.y, q) : __dadd_rn(pc.y, -q),
                                         (s_oct & 4) ? __dadd_rn(pc.z, q) : __dadd_rn(pc.z, -q), 0.5 * pc.w);
                    lk[c] = make_int4(start, ccount, sib, (L + 1) | (s_oct << 16) | (forced ? (1 << 12) : 0));
                    parent[c] = i;
                    rot[c] = rot1[rotn * 8 + pix];
                    if(topidx) topidx[c] = forced ? daughter + sl : -1;
                }
                if(sl == 0)
                    lk[i] = make_int4(tbase, nchild, link.z, wflags | (1 << 8) | (int) (mm << 24) | (forced ? (1 << 13) : 0));   // + which octants exist
            }
        }
    }
}

// Bottom-up moments of the internal nodes of one level (forcetree.cpp:1080-1100)
__global__ void __launch_bounds__(256) sq_tree_moments_kernel(double4 * __restrict__ cm, const int4 * __restrict__ lk, int ls, int le,
                                                              const double4 * __restrict__ cl, int toponly)
{
    const unsigned lane = sq_lane();
    const int sg = lane >> 3, sl = lane & 7;
    const int teams = (gridDim.x * blockDim.x) >> 3;
    const int team0 = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    for(int base_node = ls + (team0 - sg); base_node < le; base_node += teams) {
        const int i = base_node + sg;
        int4 link = make_int4(0, 0, -1, 0);
        if(i < le) link = lk[i];
        const bool internal = (i < le) && SQ_INFO_TYPE(link.w) == 1 && (!toponly || SQ_INFO_ITOP(link.w));
        double m = 0, mx = 0, my = 0, mz = 0;
        if(internal && sl < link.y) {
            const double4 c = cm[link.x + sl];
            m = c.w; mx = c.w * c.x; my = c.w * c.y; mz = c.w * c.z;
        }
        #pragma unroll
        for(int o = 4; o >= 1; o >>= 1) {
            m += __shfl_xor_sync(SQ_FULL, m, o);
            mx += __shfl_xor_sync(SQ_FULL, mx, o);
            my += __shfl_xor_sync(SQ_FULL, my, o);
            mz += __shfl_xor_sync(SQ_FULL, mz, o);
        }
        if(internal && sl == 0) {
            if(m > 0) cm[i] = make_double4(mx / m, my / m, mz / m, m);
            else if(SQ_INFO_ITOP(link.w)) { const double4 c4 = cl[i]; cm[i] = make_double4(c4.x, c4.y, c4.z, 0.0); }   // forcetree.cpp:1275-1281
            else cm[i] = make_double4(0, 0, 0, 0);
        }
    }
}

// flags[sorted position] = 1 for active particles; then compacted to an ascending list
__global__ void sq_mark_active_kernel(const int32_t * __restrict__ active, int64_t nactive, const uint32_t * __restrict__ rank, uint32_t * __restrict__ flags,
                                      int64_t n, int64_t nvalid, int * __restrict__ bad)
{
    for(int64_t k = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; k < nactive; k += (int64_t) gridDim.x * blockDim.x) {
        const int32_t a = active[k];
        if(a < 0 || a >= n) { atomicOr(bad, 1); continue; }          // an index outside the particle table
        const uint32_t r = rank[a];
        if(r < nvalid) flags[r] = 1u;                                // garbage / swallowed particles have no work (and are not in the tree)
    }
}
__global__ void sq_compact_kernel(const uint32_t * __restrict__ scanned, const uint32_t * __restrict__ flags_total, int64_t n, uint32_t * __restrict__ out)
{
    for(int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const uint32_t a = scanned[i];
        const uint32_t b = (i + 1 < n) ? scanned[i + 1] : *flags_total;
        if(b != a) out[a] = (uint32_t) i;
    }
}
// tree-order copies: out[j] = in[idx[j]]
__global__ void sq_tree_gather_kernel(const uint32_t * __restrict__ idx, int64_t m, const uint64_t * __restrict__ keys, const double * __restrict__ pos,
                                      const float * __restrict__ mass, uint64_t * __restrict__ okeys, double * __restrict__ opos, float * __restrict__ omass)
{
    for(int64_t j = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; j < m; j += (int64_t) gridDim.x * blockDim.x) {
        const int64_t s = idx[j];
        if(okeys && keys) okeys[j] = keys[s];
        opos[3 * j] = pos[3 * s]; opos[3 * j + 1] = pos[3 * s + 1]; opos[3 * j + 2] = pos[3 * s + 2];
        omass[j] = mass[s];
    }
}
__global__ void sq_topbits_kernel(const uint64_t * __restrict__ a, const uint64_t * __restrict__ b, int64_t m, int shift, int * __restrict__ flag)
{
    for(int64_t j = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; j < m; j += (int64_t) gridDim.x * blockDim.x)
        if((a[j] ^ b[j]) >> shift) atomicOr(flag, 1);
}
__global__ void sq_compose_kernel(const uint32_t * __restrict__ outer, const uint32_t * __restrict__ inner, int64_t m, uint32_t * __restrict__ out)
{
    for(int64_t j = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; j < m; j += (int64_t) gridDim.x * blockDim.x)
        out[j] = inner ? inner[outer[j]] : outer[j];
}

static int grid_for(const sq_ctx * c, int64_t n, int threads, int per_sm)
{
    int64_t need = (n + threads - 1) / threads;
    int64_t cap = (int64_t) c->num_sms * per_sm;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}


// Upload-order index list -> ascending list of sorted positions (flags -> scan -> compact).
int sq_active_to_sorted(sq_ctx * c, const int32_t * active, int64_t nactive, sq_buf * out, int64_t * count)
{
    const int64_t n = c->n;
    sq_buf & d_act = c->act_tmp;     // kept across calls: hierarchical sub-steps come by the thousand
    SQ_TRY(c->flag_tmp.reserve(sizeof(uint32_t) * (size_t) (n + 4)));
    SQ_TRY(out->reserve(sizeof(uint32_t) * (size_t) (nactive + 1)));
    SQ_TRY(d_act.reserve(sizeof(int32_t) * (size_t) (nactive + 1)));
    uint32_t * flags = c->flag_tmp.as<uint32_t>();
    int rc = SQ_OK;
    uint32_t h_total = 0;
    do {
        cudaError_t e = cudaMemcpyAsync(d_act.p, active, sizeof(int32_t) * (size_t) nactive, cudaMemcpyDefault, c->stream);   // host, managed or device list (UVA)
        if(e == cudaSuccess) e = cudaMemsetAsync(flags, 0, sizeof(uint32_t) * (size_t) (n + 4), c->stream);
        if(e != cudaSuccess) { sq_set_error("active list upload: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; break; }
        if(nactive > 0) {
            sq_mark_active_kernel<<<grid_for(c, nactive, 256, 8), 256, 0, c->stream>>>(d_act.as<int32_t>(), nactive, c->rank.as<uint32_t>(), flags, n, c->n_valid, (int *) (flags + n + 2));
            c->kernel_launches++;
        }
        uint32_t * total = flags + n + 1;
        if((rc = sq_exclusive_scan_u32(c, flags, n, total))) break;
        if(n > 0) {
            sq_compact_kernel<<<grid_for(c, n, 256, 8), 256, 0, c->stream>>>(flags, total, n, out->as<uint32_t>());
            c->kernel_launches++;
        }
        uint32_t h_two[2] = {0, 0};
        e = cudaMemcpyAsync(h_two, total, 8, cudaMemcpyDeviceToHost, c->stream);
        if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if(e != cudaSuccess) { sq_set_error("active list compaction: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; break; }
        h_total = h_two[0];
        if(h_two[1]) { sq_set_error("active list: an index lies outside [0, %lld)", (long long) n); rc = SQ_ERR_ARG; break; }
    } while(0);
    *count = h_total;
    return rc;
}

// ------------------------------------------------------------------------------------ host side
static int sq_tree_alloc(sq_ctx * c, int64_t maxnodes)
{
    SQ_TRY(c->t_cm.reserve(sizeof(double4) * (size_t) maxnodes));
    SQ_TRY(c->t_cl.reserve(sizeof(double4) * (size_t) maxnodes));
    SQ_TRY(c->t_lk.reserve(sizeof(int4) * (size_t) maxnodes));
    SQ_TRY(c->t_parent.reserve(sizeof(int) * (size_t) maxnodes));
    SQ_TRY(c->t_rot.reserve((size_t) maxnodes + 16));
    SQ_TRY(c->t_lvl.reserve(sizeof(int) * 64));
    if(c->have_domain) {
        SQ_TRY(c->t_top.reserve(sizeof(int) * (size_t) maxnodes));
        SQ_TRY(c->d_leafrange.reserve(sizeof(int2) * c->dom_leaves.size()));
    }
    return SQ_OK;
}

// Select the particle set (all, or an active subset in key order) and, if the geometric octants of
// some particle differ from its key digits, re-order the set by geometric keys.
static int sq_tree_prepare_set(sq_ctx * c, const int32_t * active, int64_t nactive)
{
    const int64_t n = c->n;
    c->tree_identity = true;
    c->tree_full = (active == nullptr);
    c->tree_np = c->n_valid;         // garbage sorts behind the valid particles and stays out of the tree (forcetree.cpp:805)
    uint32_t * order = nullptr;    // tree position -> sorted position, when not identity
    if(active) {
        int64_t h_total = 0;
        SQ_TRY(sq_active_to_sorted(c, active, nactive, &c->active_sorted, &h_total));
        c->tree_np = h_total;      // duplicates in the list collapse
        c->tree_identity = false;
        order = c->active_sorted.as<uint32_t>();
        const int64_t m = c->tree_np;
        SQ_TRY(c->tree_keys.reserve(sizeof(uint64_t) * (size_t) (m + 1)));
        SQ_TRY(c->tree_pos.reserve(sizeof(double) * 3 * (size_t) (m + 1)));
        SQ_TRY(c->tree_mass.reserve(sizeof(float) * (size_t) (m + 1)));
        SQ_TRY(c->tree_order.reserve(sizeof(uint32_t) * (size_t) (m + 1)));
        if(m > 0) {
            sq_tree_gather_kernel<<<grid_for(c, m, 256, 8), 256, 0, c->stream>>>(order, m, c->keys.as<uint64_t>(), c->pos.as<double>(), c->mass.as<float>(),
                    c->tree_keys.as<uint64_t>(), c->tree_pos.as<double>(), c->tree_mass.as<float>());
            c->kernel_launches++;
            SQ_CUDA(cudaMemcpyAsync(c->tree_order.p, order, sizeof(uint32_t) * (size_t) m, cudaMemcpyDeviceToDevice, c->stream));
        }
    }
    if(c->geo_mismatch && c->tree_np > 0) {
        // rare: order the set by geometric keys (keys kernel header comment)
        const int64_t m = c->tree_np;
        const double * spos = c->tree_identity ? c->pos.as<double>() : c->tree_pos.as<double>();
        const float * smass = c->tree_identity ? c->mass.as<float>() : c->tree_mass.as<float>();
        sq_buf gk, gv, npos, nmass, nord;
        int rc = SQ_OK;
        do {
            if((rc = gk.reserve(sizeof(uint64_t) * (size_t) m))) break;
            if((rc = gv.reserve(sizeof(uint32_t) * (size_t) m))) break;
            if((rc = npos.reserve(sizeof(double) * 3 * (size_t) m))) break;
            if((rc = nmass.reserve(sizeof(float) * (size_t) m))) break;
            if((rc = nord.reserve(sizeof(uint32_t) * (size_t) m))) break;
            if((rc = sq_geokeys_run(c, spos, m, gk.as<uint64_t>()))) break;
            if(c->have_domain && c->dom_maxlevel > 0) {
                // under a domain the re-ordering must stay inside the top leaves: a particle exactly on a face of a TOP-LEVEL cell
                // would change its top leaf (and owner); faces of the finer cells are free
                const uint64_t * pk = c->tree_identity ? c->keys.as<uint64_t>() : c->tree_keys.as<uint64_t>();
                sq_topbits_kernel<<<grid_for(c, m, 256, 8), 256, 0, c->stream>>>(gk.as<uint64_t>(), pk, m, 63 - 3 * c->dom_maxlevel, c->t_counters.as<int>() + 4);
                c->kernel_launches++;
            }
            if((rc = sq_iota(c, gv.as<uint32_t>(), m))) break;
            if((rc = sq_sort_pairs(c, gk.as<uint64_t>(), gv.as<uint32_t>(), m, 0, 64))) break;
            sq_tree_gather_kernel<<<grid_for(c, m, 256, 8), 256, 0, c->stream>>>(gv.as<uint32_t>(), m, nullptr, spos, smass, nullptr,
                    npos.as<double>(), nmass.as<float>());
            sq_compose_kernel<<<grid_for(c, m, 256, 8), 256, 0, c->stream>>>(gv.as<uint32_t>(), c->tree_identity ? nullptr : c->tree_order.as<uint32_t>(),
                    m, nord.as<uint32_t>());
            c->kernel_launches += 2;
            int h_top = 0;
            if(c->have_domain && c->dom_maxlevel > 0) cudaMemcpyAsync(&h_top, c->t_counters.as<int>() + 4, 4, cudaMemcpyDeviceToHost, c->stream);
            if(cudaStreamSynchronize(c->stream) != cudaSuccess) { sq_set_error("geometric re-order failed"); rc = SQ_ERR_CUDA; break; }
            if(h_top) { sq_set_error("tree build: a particle lies exactly on a face of a top-level cell of the domain"); rc = SQ_ERR_ARG; break; }
        } while(0);
        if(rc != SQ_OK) { gk.release(); gv.release(); npos.release(); nmass.release(); nord.release(); return rc; }
        // adopt the re-ordered buffers
        c->tree_keys.release(); c->tree_pos.release(); c->tree_mass.release(); c->tree_order.release();
        c->tree_keys = gk; c->tree_pos = npos; c->tree_mass = nmass; c->tree_order = nord;
        gv.release();
        c->tree_identity = false;
    }
    return SQ_OK;
}

int sq_tree_run(sq_ctx * c, const int32_t * active, int64_t nactive, sq_tree_info * info)
{
    if(!c->sorted) { sq_set_error("sq_tree_build: call sq_ph_sort first"); return SQ_ERR_ARG; }
    c->tree_built = false;
    SQ_CUDA(cudaEventRecord(c->ev[4], c->stream));
    SQ_TRY(sq_tree_prepare_set(c, active, nactive));
    const int64_t np = c->tree_np;
    const uint64_t * tk = c->tree_identity ? c->keys.as<uint64_t>() : c->tree_keys.as<uint64_t>();
    const double * tp = c->tree_identity ? c->pos.as<double>() : c->tree_pos.as<double>();
    const float * tmass = c->tree_identity ? c->mass.as<float>() : c->tree_mass.as<float>();
    double factor = c->par.TreeAllocFactor > 0 ? c->par.TreeAllocFactor : 0.9;
    const bool dom = c->have_domain;
    const sq_topnode * d_tn = dom ? c->d_topnodes.as<sq_topnode>() : nullptr;
    const sq_topleaf * d_tl = dom ? c->d_topleaves.as<sq_topleaf>() : nullptr;
    const int ntopnodes = dom ? (int) c->dom_nodes.size() : 0;
    const unsigned char * phtab = sq_phtab_device();
    if(!phtab) return SQ_ERR_CUDA;
    int h_lvl[SQ_MAXLEVEL + 4];
    int h_cnt[4];
    for(;;) {
        // forcetree.cpp:200: maxnodes = TreeAllocFactor * NumPart + NTopNodes
        int64_t maxnodes = (int64_t) (factor * (double) np) + 1 + 64 + ntopnodes;
        if(maxnodes > 0x5fffffff) maxnodes = 0x5fffffff;   // forcetree.cpp:1385 index limit
        SQ_TRY(sq_tree_alloc(c, maxnodes));
        int * lvl = c->t_lvl.as<int>();
        int * cnt = c->t_counters.as<int>();
        SQ_CUDA(cudaMemsetAsync(lvl, 0, sizeof(int) * 64, c->stream));
        sq_tree_root_kernel<<<1, 1, 0, c->stream>>>(c->t_cm.as<double4>(), c->t_cl.as<double4>(), c->t_lk.as<int4>(), c->t_parent.as<int>(),
                c->t_rot.as<unsigned char>(), lvl, cnt, (int) np, c->par.BoxSize, dom ? c->t_top.as<int>() : nullptr);
        if(dom) SQ_CUDA(cudaMemsetAsync(c->d_treenode.p, 0xff, sizeof(int) * c->dom_leaves.size(), c->stream));
        c->kernel_launches++;
        const int grid = c->num_sms * 4;
        for(int L = 0; L <= SQ_MAXLEVEL; L++) {
            sq_tree_split_kernel<<<grid, 256, 0, c->stream>>>(tk, tp, tmass, c->t_cm.as<double4>(), c->t_cl.as<double4>(), c->t_lk.as<int4>(),
                    c->t_parent.as<int>(), c->t_rot.as<unsigned char>(), lvl, cnt, L, (int) maxnodes, phtab,
                    d_tn, d_tl, dom ? c->d_treenode.as<int>() : nullptr, dom ? c->t_top.as<int>() : nullptr, c->comm_rank, dom ? c->d_leafrange.as<int2>() : nullptr);
            sq_tree_advance_kernel<<<1, 1, 0, c->stream>>>(lvl, cnt, L, (int) maxnodes);
            c->kernel_launches += 2;
        }
        SQ_CUDA(cudaGetLastError());
        SQ_CUDA(cudaMemcpyAsync(h_lvl, lvl, sizeof(int) * (SQ_MAXLEVEL + 3), cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaMemcpyAsync(h_cnt, cnt, sizeof(int) * 4, cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        if(h_cnt[CNT_ERR] & 4) { sq_set_error("tree build: a local particle lies in a top leaf that the domain assigns to another rank (exchange particles first)"); return SQ_ERR_ARG; }
        if(h_cnt[CNT_ERR] & 2) { sq_set_error("tree build: more than 8 particles share one 2^-21 cell"); return SQ_ERR_TREE_DEPTH; }
        if(h_cnt[CNT_ERR] & 1) {
            // forcetree.cpp:215-231: grow the pool by 1.15, give up above 3.0
            factor *= 1.15;
            if(factor > 3.0 && np > 64) { sq_set_error("tree build: TreeAllocFactor %g is too large", factor); return SQ_ERR_NODE_POOL; }
            continue;
        }
        c->tree_maxnodes = maxnodes;
        c->tree_numnodes = h_cnt[CNT_NEXT];
        break;
    }
    if(c->par.TreeAllocFactor > 0) c->par.TreeAllocFactor = factor;
    int maxlevel = 0;
    for(int L = 0; L <= SQ_MAXLEVEL; L++) {
        c->lvl_start[L] = h_lvl[L];
        if(h_lvl[L + 1] > h_lvl[L]) maxlevel = L;
    }
    c->lvl_start[SQ_MAXLEVEL + 1] = h_lvl[SQ_MAXLEVEL + 1];
    c->tree_maxlevel = maxlevel;
    SQ_CUDA(cudaEventRecord(c->ev[5], c->stream));
    for(int L = maxlevel - 1; L >= 0; L--) {
        const int ls = h_lvl[L], le = h_lvl[L + 1];
        const int g = grid_for(c, (int64_t) (le - ls) * 8, 256, 8);
        sq_tree_moments_kernel<<<g, 256, 0, c->stream>>>(c->t_cm.as<double4>(), c->t_lk.as<int4>(), ls, le, c->t_cl.as<double4>(), 0);
        c->kernel_launches++;
    }
    SQ_CUDA(cudaGetLastError());
    SQ_CUDA(cudaEventRecord(c->ev[6], c->stream));
    double4 h_root = make_double4(0, 0, 0, 0);
    SQ_CUDA(cudaMemcpyAsync(&h_root, c->t_cm.p, sizeof(h_root), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    c->tree_total_mass = h_root.w;
    c->n_local_nodes = c->tree_numnodes;
    c->n_sources = np;
    c->let_attached = false;
    c->walk_munit = 0;
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]); c->tm.tree_build_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[5], c->ev[6]); c->tm.tree_moments_ms = ms;
    c->tree_built = true;
    if(info) {
        info->numnodes = c->tree_numnodes;
        info->numparticles = np;
        info->maxnodes = c->tree_maxnodes;
        info->maxlevel = maxlevel;
        info->geo_mismatch = c->geo_mismatch;
    }
    return SQ_OK;
}

// internal top-level nodes of one level from their 8 children (force_treeupdate_pseudos, forcetree.cpp:1212-1282)
int sq_tree_moments_top(sq_ctx * c, int ls, int le)
{
    if(le <= ls) return SQ_OK;
    const int g = grid_for(c, (int64_t) (le - ls) * 8, 256, 8);
    sq_tree_moments_kernel<<<g, 256, 0, c->stream>>>(c->t_cm.as<double4>(), c->t_lk.as<int4>(), ls, le, c->t_cl.as<double4>(), 1);
    c->kernel_launches++;
    return SQ_OK;
}

// ------------------------------------------------------------------------------------ NODE image
// struct NODE, libgadget/forcetree.h:38-66 (120 bytes)
struct sq_ref_node {
    int sibling, father;
    double len, center[3], cofm[3], mass, hmax;
    int suns[8];
    int noccupied;
    unsigned int flags;
};
static_assert(sizeof(sq_ref_node) == 120, "NODE image must be 120 bytes");

__global__ void sq_tree_export_kernel(const double4 * __restrict__ cm, const double4 * __restrict__ cl, const int4 * __restrict__ lk,
                                      const int * __restrict__ parent, int numnodes, long long firstnode,
                                      const uint32_t * __restrict__ tree_order, const uint32_t * __restrict__ perm, sq_ref_node * __restrict__ out,
                                      const int * __restrict__ topidx, const sq_topnode * __restrict__ tn, const sq_topleaf * __restrict__ tl, int myrank,
                                      long long lastnode)
{
    for(int i = blockIdx.x * blockDim.x + threadIdx.x; i < numnodes; i += gridDim.x * blockDim.x) {
        int4 l = lk[i];
        const double4 c = cl[i], m = cm[i];
        // a top leaf of another rank is a pseudo node in the reference's image, whatever was imported below it on the device:
        // ChildType PSEUDO, suns[0] = lastnode + leaf (force_create_node_for_topnode, forcetree.cpp:903-906)
        int pseudo_leaf = -1;
        if(topidx && topidx[i] >= 0) {
            const sq_topnode T = tn[topidx[i]];
            if(T.Daughter < 0 && tl[T.Leaf].Task != myrank) pseudo_leaf = T.Leaf;
        }
        if(pseudo_leaf >= 0) l = make_int4(0, 0, l.z, (l.w & ~(3 << 8)) | (2 << 8));
        sq_ref_node nd;
        nd.sibling = l.z >= 0 ? (int) (firstnode + l.z) : -1;
        nd.father = parent[i] >= 0 ? (int) (firstnode + parent[i]) : -1;
        nd.len = c.w; nd.center[0] = c.x; nd.center[1] = c.y; nd.center[2] = c.z;
        nd.cofm[0] = m.x; nd.cofm[1] = m.y; nd.cofm[2] = m.z; nd.mass = m.w; nd.hmax = 0;
        const int type = SQ_INFO_TYPE(l.w);
        for(int j = 0; j < 8; j++) {
            int v = -1;
            if(j < l.y) {
                if(type == 1) v = (int) (firstnode + l.x + j);
                else {
                    uint32_t p = (uint32_t) (l.x + j);
                    if(tree_order) p = tree_order[p];
                    if(perm) p = perm[p];
                    v = (int) p;
                }
            }
            nd.suns[j] = v;
        }
        if(pseudo_leaf >= 0) nd.suns[0] = (int) (lastnode + pseudo_leaf);
        nd.noccupied = type == 1 ? SQ_NODEFULL : l.y;
        nd.flags = (unsigned) SQ_INFO_ITOP(l.w) | ((unsigned) SQ_INFO_TOP(l.w) << 1) | ((unsigned) SQ_INFO_LOCAL(l.w) << 2) | ((unsigned) type << 3);
        out[i] = nd;
    }
}

int sq_tree_export(sq_ctx * c, void * host_nodes, int64_t firstnode, int upload_order, int64_t lastnode)
{
    if(!c->tree_built) { sq_set_error("sq_tree_export_nodes: no tree"); return SQ_ERR_ARG; }
    const int nn = (int) c->n_local_nodes;          // imported sub-trees are not part of the host image
    if(lastnode < 0) lastnode = firstnode + nn;
    sq_buf img;
    SQ_TRY(img.reserve(sizeof(sq_ref_node) * (size_t) nn));
    sq_tree_export_kernel<<<grid_for(c, nn, 256, 8), 256, 0, c->stream>>>(c->t_cm.as<double4>(), c->t_cl.as<double4>(), c->t_lk.as<int4>(),
            c->t_parent.as<int>(), nn, (long long) firstnode, c->tree_identity ? nullptr : c->tree_order.as<uint32_t>(),
            upload_order ? c->perm.as<uint32_t>() : nullptr, img.as<sq_ref_node>(),
            c->have_domain ? c->t_top.as<int>() : nullptr, c->have_domain ? c->d_topnodes.as<sq_topnode>() : nullptr,
            c->have_domain ? c->d_topleaves.as<sq_topleaf>() : nullptr, c->comm_rank, (long long) lastnode);
    c->kernel_launches++;
    cudaError_t e = cudaMemcpyAsync(host_nodes, img.p, sizeof(sq_ref_node) * (size_t) nn, cudaMemcpyDeviceToHost, c->stream);
    if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    img.release();
    if(e != cudaSuccess) { sq_set_error("sq_tree_export_nodes: %s", cudaGetErrorString(e)); return SQ_ERR_CUDA; }
    return SQ_OK;
}
