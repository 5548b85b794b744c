// Short-range tree walk: the interaction kernel.
//
// Replaces GravLocalTreeWalk::visit<PRIMARY> / apply_accn / shall_we_discard_node / shall_we_open_node
// (libgadget/gravshort2.hpp:152-358), TreeWalk::ev_primary (treewalk2.h:566-595), the GPU kernels
// treewalk_primary_kernel / treewalk_postprocess_kernel (treewalk2.cuh:104-173) and
// GravTreeOutput::postprocess (gravshort2.hpp:88-107).
//
// Design (not the reference's thread-per-particle depth-first walk):
//   * One warp walks the tree for a GROUP of G Peano-Hilbert-adjacent active particles (G = 32, 16, 8;
//     the 32/G lane slices of a target split the interaction list between them).
//   * The warp shares ONE node stack and ONE interaction list, both in shared memory.  A stack entry is
//     (node, mask): mask = the targets of the group that reach this node, i.e. opened every ancestor.
//     Per step the warp pops up to 32 entries, one per lane; node records come in as coalesced 128-bit
//     loads (children of a node are contiguous).
//   * Each lane applies the reference's per-particle tests (discard: r2 vs Rcut and the Rcut+len/2
//     box; open: relative-acceleration criterion M l^2 > r^4 alpha |a_old|, Barnes-Hut angle,
//     inside-0.6-len box) to ITS node for every target in the mask, in FP64 with the reference's
//     expressions, so every (particle, node) decision is the one the reference takes.  That yields an
//     accept mask (targets that use the node's monopole) and an open mask (targets that descend).
//     Accepted nodes and the particles of opened leaves go to the interaction list together with their
//     masks; children are pushed with the open mask.  A group-level bounding-box test discards nodes
//     no target can reach without running the per-target loop.
//   * The list stores sources as FP32 offsets from the group centre (formed in FP64, periodic wrap
//     applied there, once per source rather than once per pair).  Evaluation is FP32: rsqrt, spline
//     softening, the erfc/"exact" split factor from a shared-memory-resident table (value, slope pairs
//     -> one 64-bit LDS + one FMA), FMA accumulation.  Partial sums are carried in FP32 for at most one
//     list flush (<= LIST_CAP entries) and then added to FP64 accumulators.
//   * Epilogue = postprocess: x G, potential self terms; results written in sorted order.
#include "sq_common.cuh"
#include "shortrange_table.h"
#include <math.h>

#define WALK_WARPS 8
#define WALK_THREADS (WALK_WARPS * 32)
#define STACK_CAP 512
#define STACK_RESERVE 160   /* 7 siblings x 22 levels: room a depth-first descent can always need */
#define LIST_CAP 128

struct sq_walk_par {
    double box, half, rcut, rcut2, bh2, h;          // h = ForceSoftening = 2.8 eps
    double G, cbrtrho0, errtol;
    int relative;                                   // TreeUseBH == 0
    float f_box, f_invbox, f_h2, f_inv_cell_dx, f_range;   // f_range: distance beyond which the table gives zero
};

struct sq_walk_warp_smem {
    double tx[32], ty[32], tz[32], ta[32];
    int2 stack[STACK_CAP];
    float4 lst[LIST_CAP];
    unsigned lmask[LIST_CAP];
    int s_end[32], s_first[32];
    unsigned s_mask[32];
};

// spline-softened force and potential factors, gravshort2.hpp:332-350 (evaluated in FP64: rare branch)
__device__ __noinline__ void sq_soft_kernel(float r, float m, double h, float * fac, float * facpot)
{
    const double h3_inv = 1.0 / h / h / h;
    const double u = (double) r / h;
    double f, wp;
    if(u < 0.5) {
        f = (double) m * h3_inv * (10.666666666667 + u * u * (32.0 * u - 38.4));
        wp = -2.8 + u * u * (5.333333333333 + u * u * (6.4 * u - 9.6));
    } else {
        f = (double) m * h3_inv * (21.333333333333 - 48.0 * u + 38.4 * u * u - 10.666666666667 * u * u * u - 0.066666666667 / (u * u * u));
        wp = -3.2 + 0.066666666667 / u + u * u * (10.666666666667 + u * (-16.0 + u * (9.6 - 2.133333333333 * u)));
    }
    *fac = (float) f;
    *facpot = (float) ((double) m / h * wp);
}

template <int G, bool POT>
struct sq_walk_acc {
    float ax, ay, az, ap;
    double dax, day, daz, dap;
    unsigned nint;
};

// Evaluate list entries [0, n) for this lane's target (slice sl of S takes entries sl, sl+S, ...).
template <int G, bool POT>
__device__ __forceinline__ void sq_walk_flush(sq_walk_warp_smem & W, int n, const float2 * __restrict__ ftab, const float2 * __restrict__ ptab,
                                              const sq_walk_par & P, float txr, float tyr, float tzr, int tl, int sl, bool pairwrap,
                                              sq_walk_acc<G, POT> & A)
{
    constexpr int S = 32 / G;
    for(int e = sl; e < n; e += S) {
        const float4 s = W.lst[e];
        const unsigned m = W.lmask[e];
        if((m >> tl) & 1u) {
            float dx = s.x - txr, dy = s.y - tyr, dz = s.z - tzr;
            if(pairwrap) {
                dx -= P.f_box * rintf(dx * P.f_invbox);
                dy -= P.f_box * rintf(dy * P.f_invbox);
                dz -= P.f_box * rintf(dz * P.f_invbox);
            }
            const float r2 = dx * dx + dy * dy + dz * dz;
            const float rinv = rsqrtf(r2);
            float r = r2 * rinv;
            float fac = s.w * rinv * rinv * rinv;
            float facpot = -s.w * rinv;
            if(r2 < P.f_h2) {
                r = sqrtf(r2);
                sq_soft_kernel(r, s.w, P.h, &fac, &facpot);
            }
            A.nint++;
            // gravity.h:48-60: i = r/cellsize/dx, zero beyond the table, linear interpolation
            const float x = r * P.f_inv_cell_dx;
            if(x < (float) (SQ_NTAB - 1)) {
                const float t = (x - 0.5f) + 12582912.f;                 // round(x - 0.5) in the mantissa
                const int idx = __float_as_int(t) - 0x4B400000;
                const float frac = x - (t - 12582912.f);
                const float2 fw = ftab[idx];
                fac *= fmaf(frac, fw.y, fw.x);
                A.ax = fmaf(dx, fac, A.ax);
                A.ay = fmaf(dy, fac, A.ay);
                A.az = fmaf(dz, fac, A.az);
                if(POT) {
                    const float2 pw = ptab[idx];
                    A.ap = fmaf(facpot, fmaf(frac, pw.y, pw.x), A.ap);
                }
            }
        }
    }
    A.dax += (double) A.ax; A.day += (double) A.ay; A.daz += (double) A.az;
    A.ax = A.ay = A.az = 0.f;
    if(POT) { A.dap += (double) A.ap; A.ap = 0.f; }
}

template <int G, bool POT>
__global__ void __launch_bounds__(WALK_THREADS, 2)
sq_walk_kernel(const double4 * __restrict__ cm, const double4 * __restrict__ cl, const int4 * __restrict__ lk,
               const double * __restrict__ spos, const float * __restrict__ smass,
               const double * __restrict__ tpos, const float * __restrict__ tmass, const double * __restrict__ toldacc,
               const uint32_t * __restrict__ targets, long long ntargets, const float2 * __restrict__ gtab,
               double * __restrict__ accel, double * __restrict__ pot, unsigned long long * __restrict__ stats,
               int * __restrict__ group_counter, const sq_walk_par P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2 * ftab = reinterpret_cast<float2 *>(smem_raw);
    float2 * ptab = ftab + SQ_NTAB;
    sq_walk_warp_smem * Wall = reinterpret_cast<sq_walk_warp_smem *>(smem_raw + 2 * SQ_NTAB * sizeof(float2));
    for(int i = threadIdx.x; i < 2 * SQ_NTAB; i += blockDim.x)
        ftab[i] = gtab[i];
    __syncthreads();

    constexpr int S = 32 / G;
    const unsigned lane = sq_lane();
    const int tl = lane % G, sl = lane / G;
    sq_walk_warp_smem & W = Wall[threadIdx.x >> 5];
    const long long ngroups = (ntargets + G - 1) / G;
    unsigned long long st_int = 0, st_ent = 0, st_tst = 0;

    for(;;) {
        long long g = 0;
        if(lane == 0) g = atomicAdd(group_counter, 1);
        g = __shfl_sync(SQ_FULL, g, 0);
        if(g >= ngroups) break;
        const long long t0 = g * G;
        const int nt = (int) ((ntargets - t0) < G ? (ntargets - t0) : G);
        const bool valid = tl < nt;
        long long tidx = 0;
        double px = 0, py = 0, pz = 0, aold = 0;
        if(valid) {
            tidx = targets ? (long long) targets[t0 + tl] : (t0 + tl);
            px = tpos[3 * tidx]; py = tpos[3 * tidx + 1]; pz = tpos[3 * tidx + 2];
            aold = P.errtol * toldacc[tidx];
        }
        // group bounding box over the valid targets
        double lox = valid ? px : 1e300, loy = valid ? py : 1e300, loz = valid ? pz : 1e300;
        double hix = valid ? px : -1e300, hiy = valid ? py : -1e300, hiz = valid ? pz : -1e300;
        #pragma unroll
        for(int o = 16; o >= 1; o >>= 1) {
            lox = fmin(lox, __shfl_xor_sync(SQ_FULL, lox, o)); hix = fmax(hix, __shfl_xor_sync(SQ_FULL, hix, o));
            loy = fmin(loy, __shfl_xor_sync(SQ_FULL, loy, o)); hiy = fmax(hiy, __shfl_xor_sync(SQ_FULL, hiy, o));
            loz = fmin(loz, __shfl_xor_sync(SQ_FULL, loz, o)); hiz = fmax(hiz, __shfl_xor_sync(SQ_FULL, hiz, o));
        }
        const double gcx = 0.5 * (lox + hix), gcy = 0.5 * (loy + hiy), gcz = 0.5 * (loz + hiz);
        const double ghx = 0.5 * (hix - lox), ghy = 0.5 * (hiy - loy), ghz = 0.5 * (hiz - loz);
        __syncwarp();
        if(sl == 0) {
            W.tx[tl] = px; W.ty[tl] = py; W.tz[tl] = pz; W.ta[tl] = aold;
        }
        const float txr = (float) (px - gcx), tyr = (float) (py - gcy), tzr = (float) (pz - gcz);
        // pairs may straddle the periodic image once group extent + table range reaches half the box
        const bool pairwrap = 2.0 * fmax(ghx, fmax(ghy, ghz)) + (double) P.f_range >= P.half;
        sq_walk_acc<G, POT> A;
        A.ax = A.ay = A.az = A.ap = 0.f;
        A.dax = A.day = A.daz = A.dap = 0.0;
        A.nint = 0;
        int top = 0, nlist = 0;
        if(lane == 0) W.stack[0] = make_int2(0, nt >= 32 ? (int) 0xffffffffu : (int) ((1u << nt) - 1u));
        top = 1;
        __syncwarp();

        while(top > 0) {
            // pop: as many as fit while keeping room for 8 children each
            // wide pops only while a depth-first descent (<= 7 entries per level) still fits above them
            int k = (STACK_CAP - STACK_RESERVE - top) / 7;
            k = k < 1 ? 1 : k;
            k = k > 32 ? 32 : k;
            k = k > top ? top : k;
            top -= k;
            const bool have = (int) lane < k;
            int node = 0;
            unsigned mask = 0;
            if(have) {
                const int2 e = W.stack[top + lane];
                node = e.x; mask = (unsigned) e.y;
            }
            __syncwarp();
            unsigned acc_mask = 0, open_mask = 0;
            int4 L = make_int4(0, 0, -1, 0);
            double4 M = make_double4(0, 0, 0, 0);
            if(have) {
                L = lk[node];
                M = cm[node];
                const double4 C = cl[node];
                const double len = C.w;
                st_tst++;
                // group-level discard (gravshort2.hpp:152-167 holds for every target if it holds for the box)
                const double eff = P.rcut + 0.5 * len;
                const double bx = fabs(sq_nearest(C.x - gcx, P.box, P.half)) - ghx;
                const double by = fabs(sq_nearest(C.y - gcy, P.box, P.half)) - ghy;
                const double bz = fabs(sq_nearest(C.z - gcz, P.box, P.half)) - ghz;
                if(!(bx > eff || by > eff || bz > eff)) {
                    const double inside = 0.6 * len;
                    const double ml2 = M.w * len * len;
                    const double l2 = len * len;
                    unsigned mm = mask;
                    while(mm) {
                        const int t = __ffs(mm) - 1;
                        mm &= mm - 1;
                        const double x = W.tx[t], y = W.ty[t], z = W.tz[t];
                        const double dx = sq_nearest(M.x - x, P.box, P.half);
                        const double dy = sq_nearest(M.y - y, P.box, P.half);
                        const double dz = sq_nearest(M.z - z, P.box, P.half);
                        const double r2 = dx * dx + dy * dy + dz * dz;
                        const double ex = fabs(sq_nearest(C.x - x, P.box, P.half));
                        const double ey = fabs(sq_nearest(C.y - y, P.box, P.half));
                        const double ez = fabs(sq_nearest(C.z - z, P.box, P.half));
                        // shall_we_discard_node, gravshort2.hpp:152-167
                        if(r2 > P.rcut2 && (ex > eff || ey > eff || ez > eff))
                            continue;
                        // shall_we_open_node, gravshort2.hpp:172-193
                        bool open = false;
                        if(P.relative && (ml2 > r2 * r2 * W.ta[t])) open = true;
                        else if(l2 / r2 > P.bh2) open = true;
                        else if(ex < inside && ey < inside && ez < inside) open = true;
                        if(open) open_mask |= 1u << t;
                        else acc_mask |= 1u << t;
                    }
                }
            }
            const int type = SQ_INFO_TYPE(L.w);
            // ---- accepted nodes -> interaction list
            {
                const unsigned b = __ballot_sync(SQ_FULL, acc_mask != 0);
                const int tot = __popc(b);
                if(tot) {
                    if(nlist + tot > LIST_CAP) {
                        __syncwarp();
                        st_ent += nlist;
                        sq_walk_flush<G, POT>(W, nlist, ftab, ptab, P, txr, tyr, tzr, tl, sl, pairwrap, A);
                        nlist = 0;
                        __syncwarp();
                    }
                    if(acc_mask) {
                        const int p = nlist + __popc(b & sq_lanemask_lt());
                        W.lst[p] = make_float4((float) sq_nearest(M.x - gcx, P.box, P.half), (float) sq_nearest(M.y - gcy, P.box, P.half),
                                               (float) sq_nearest(M.z - gcz, P.box, P.half), (float) M.w);
                        W.lmask[p] = acc_mask;
                    }
                    nlist += tot;
                }
            }
            // ---- opened internal nodes -> push children with the open mask (pseudo nodes are skipped: gravshort2.hpp:305-314)
            {
                const int nch = (open_mask && type == 1) ? L.y : 0;
                int tot;
                const int off = sq_warp_excl_scan(nch, tot);
                for(int j = 0; j < nch; j++)
                    W.stack[top + off + j] = make_int2(L.x + j, (int) open_mask);
                top += tot;
            }
            // ---- opened leaves -> their particles join the list (gravshort2.hpp:291-303)
            {
                const int np = (open_mask && type == 0) ? L.y : 0;
                int tot;
                const int off = sq_warp_excl_scan(np, tot);
                if(tot) {
                    W.s_end[lane] = off + np;
                    W.s_first[lane] = L.x - off;        // particle = s_first + j
                    W.s_mask[lane] = open_mask;
                    __syncwarp();
                    for(int base = 0; base < tot; base += 32) {
                        const int chunk = (tot - base) < 32 ? (tot - base) : 32;
                        if(nlist + chunk > LIST_CAP) {
                            __syncwarp();
                            st_ent += nlist;
                            sq_walk_flush<G, POT>(W, nlist, ftab, ptab, P, txr, tyr, tzr, tl, sl, pairwrap, A);
                            nlist = 0;
                            __syncwarp();
                        }
                        const int j = base + lane;
                        if(j < tot) {
                            // first source lane whose end is past j
                            int lo = 0, hi = 31;
                            while(lo < hi) {
                                const int mid = (lo + hi) >> 1;
                                if(W.s_end[mid] > j) hi = mid; else lo = mid + 1;
                            }
                            const long long p = (long long) W.s_first[lo] + j;
                            const double sx = spos[3 * p], sy = spos[3 * p + 1], sz = spos[3 * p + 2];
                            W.lst[nlist + lane] = make_float4((float) sq_nearest(sx - gcx, P.box, P.half), (float) sq_nearest(sy - gcy, P.box, P.half),
                                                              (float) sq_nearest(sz - gcz, P.box, P.half), smass[p]);
                            W.lmask[nlist + lane] = W.s_mask[lo];
                        }
                        nlist += chunk;
                    }
                    __syncwarp();
                }
            }
            __syncwarp();
        }
        __syncwarp();
        st_ent += nlist;
        sq_walk_flush<G, POT>(W, nlist, ftab, ptab, P, txr, tyr, tzr, tl, sl, pairwrap, A);
        // combine the slices of each target
        #pragma unroll
        for(int o = G; o < 32; o <<= 1) {
            A.dax += __shfl_xor_sync(SQ_FULL, A.dax, o);
            A.day += __shfl_xor_sync(SQ_FULL, A.day, o);
            A.daz += __shfl_xor_sync(SQ_FULL, A.daz, o);
            if(POT) A.dap += __shfl_xor_sync(SQ_FULL, A.dap, o);
            A.nint += __shfl_xor_sync(SQ_FULL, A.nint, o);
        }
        if(valid && sl == 0) {
            // gravshort2.hpp:88-107
            accel[3 * tidx] = A.dax * P.G;
            accel[3 * tidx + 1] = A.day * P.G;
            accel[3 * tidx + 2] = A.daz * P.G;
            if(POT) {
                const double m = (double) tmass[tidx];
                double p = A.dap;
                p += m / (P.h / 2.8);
                p -= 2.8372975 * pow(m, 2.0 / 3) * P.cbrtrho0;
                pot[tidx] = p * P.G;
            }
            st_int += A.nint;
        }
        __syncwarp();
    }
    // per-warp statistics
    #pragma unroll
    for(int o = 16; o >= 1; o >>= 1) {
        st_int += __shfl_xor_sync(SQ_FULL, st_int, o);
        st_tst += __shfl_xor_sync(SQ_FULL, st_tst, o);
    }
    if(lane == 0) {
        atomicAdd(&stats[0], st_int);
        atomicAdd(&stats[1], st_ent);
        atomicAdd(&stats[2], st_tst);
    }
}

// ------------------------------------------------------------------------------------ host side
static const double h_sr_x[SQ_NGRAVTAB] = SQ_SHORTRANGE_X_INIT;
static const double h_sr_pot[SQ_NGRAVTAB] = SQ_SHORTRANGE_POT_INIT;
static const double h_sr_force[SQ_NGRAVTAB] = SQ_SHORTRANGE_FORCE_INIT;

// GravShortTable, libgadget/gravshort-tree2.cpp:29-69: float tables; here pre-differenced (value, slope)
static int sq_walk_table(sq_ctx * c)
{
    float ft[SQ_NTAB], pt[SQ_NTAB];
    for(int i = 0; i < SQ_NTAB; i++) {
        const double u = h_sr_x[i] * 0.5 / c->par.Asmth;
        if(c->par.WindowType == 1) {
            ft[i] = (float) h_sr_force[i];
            pt[i] = (float) h_sr_pot[i];
        } else {
            ft[i] = (float) (erfc(u) + 2.0 * u / sqrt(M_PI) * exp(-u * u));
            pt[i] = (float) erfc(u);
        }
    }
    float2 h[2 * SQ_NTAB];
    for(int i = 0; i < SQ_NTAB; i++) {
        const int j = i + 1 < SQ_NTAB ? i + 1 : i;
        h[i] = make_float2(ft[i], (float) ((double) ft[j] - (double) ft[i]));
        h[SQ_NTAB + i] = make_float2(pt[i], (float) ((double) pt[j] - (double) pt[i]));
    }
    SQ_TRY(c->w_table.reserve(sizeof(h)));
    SQ_CUDA(cudaMemcpyAsync(c->w_table.p, h, sizeof(h), cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    return SQ_OK;
}

template <int G, bool POT>
static int sq_walk_launch(sq_ctx * c, const uint32_t * targets, int64_t ntargets, const sq_walk_par & P)
{
    const size_t smem = 2 * SQ_NTAB * sizeof(float2) + WALK_WARPS * sizeof(sq_walk_warp_smem);
    SQ_CUDA(cudaFuncSetAttribute(sq_walk_kernel<G, POT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    int per_sm = 0;
    SQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sq_walk_kernel<G, POT>, WALK_THREADS, smem));
    if(per_sm < 1) per_sm = 1;
    const int64_t ngroups = (ntargets + G - 1) / G;
    int64_t grid = (int64_t) c->num_sms * per_sm;
    const int64_t need = (ngroups + WALK_WARPS - 1) / WALK_WARPS;
    if(grid > need) grid = need;
    if(grid < 1) grid = 1;
    const double * spos = c->tree_identity ? c->pos.as<double>() : c->tree_pos.as<double>();
    const float * smass = c->tree_identity ? c->mass.as<float>() : c->tree_mass.as<float>();
    sq_walk_kernel<G, POT><<<(unsigned) grid, WALK_THREADS, smem, c->stream>>>(c->t_cm.as<double4>(), c->t_cl.as<double4>(), c->t_lk.as<int4>(),
            spos, smass, c->pos.as<double>(), c->mass.as<float>(), c->oldacc.as<double>(), targets, (long long) ntargets,
            c->w_table.as<float2>(), c->w_accel.as<double>(), c->w_pot.as<double>(), c->w_stats.as<unsigned long long>(),
            c->w_counter.as<int>(), P);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

int sq_walk_group_width = 32;   // tunable: targets per warp (32, 16 or 8)

int sq_walk_run(sq_ctx * c, const int32_t * active, int64_t nactive, sq_walk_stats * stats)
{
    if(!c->tree_built) { sq_set_error("sq_tree_walk: build the tree first (moments not computed)"); return SQ_ERR_ARG; }
    const sq_params & par = c->par;
    SQ_TRY(sq_walk_table(c));
    sq_walk_par P;
    memset(&P, 0, sizeof(P));
    // GravTreeParams, libgadget/gravshort2.hpp:45-54
    const double cellsize = par.BoxSize / par.Nmesh;
    P.box = par.BoxSize;
    P.half = 0.5 * par.BoxSize;
    P.rcut = par.TreeRcut * par.Asmth * cellsize;
    P.rcut2 = P.rcut * P.rcut;
    P.h = 2.8 * par.GravitySoftening;
    P.G = par.G;
    P.cbrtrho0 = pow(par.rho0, 1.0 / 3);
    P.errtol = par.ErrTolForceAcc;
    P.relative = par.TreeUseBH == 0;
    P.bh2 = par.BHOpeningAngle * par.BHOpeningAngle;
    if(par.TreeUseBH == 0) P.bh2 = par.MaxBHOpeningAngle * par.MaxBHOpeningAngle;
    P.f_box = (float) par.BoxSize;
    P.f_invbox = (float) (1.0 / par.BoxSize);
    P.f_h2 = (float) (P.h * P.h);
    P.f_inv_cell_dx = (float) (1.0 / (cellsize * SQ_SHORTRANGE_DX));
    P.f_range = (float) ((SQ_NTAB - 1) * cellsize * SQ_SHORTRANGE_DX);

    const int64_t n = c->n;
    const uint32_t * targets = nullptr;
    int64_t ntargets = n;
    if(active) {
        SQ_TRY(sq_active_to_sorted(c, active, nactive, &c->w_targets, &ntargets));
        targets = c->w_targets.as<uint32_t>();
    }
    const bool pot = c->tree_full;     // update_potential = tree->full_particle_tree_flag (gravshort-tree2.cpp:148)
    SQ_TRY(c->w_accel.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
    SQ_TRY(c->w_pot.reserve(sizeof(double) * (size_t) (n + 1)));
    SQ_TRY(c->w_stats.reserve(64));
    SQ_TRY(c->w_counter.reserve(64));
    SQ_CUDA(cudaMemsetAsync(c->w_stats.p, 0, 64, c->stream));
    SQ_CUDA(cudaMemsetAsync(c->w_counter.p, 0, 64, c->stream));
    SQ_CUDA(cudaMemsetAsync(c->w_accel.p, 0, sizeof(double) * 3 * (size_t) n, c->stream));
    if(pot) SQ_CUDA(cudaMemsetAsync(c->w_pot.p, 0, sizeof(double) * (size_t) n, c->stream));
    SQ_CUDA(cudaEventRecord(c->ev[8], c->stream));
    const int G = sq_walk_group_width;
    if(ntargets > 0) {
        int rc;
        if(G == 8) rc = pot ? sq_walk_launch<8, true>(c, targets, ntargets, P) : sq_walk_launch<8, false>(c, targets, ntargets, P);
        else if(G == 16) rc = pot ? sq_walk_launch<16, true>(c, targets, ntargets, P) : sq_walk_launch<16, false>(c, targets, ntargets, P);
        else rc = pot ? sq_walk_launch<32, true>(c, targets, ntargets, P) : sq_walk_launch<32, false>(c, targets, ntargets, P);
        SQ_TRY(rc);
    }
    SQ_CUDA(cudaEventRecord(c->ev[9], c->stream));
    unsigned long long h_stats[4] = {0, 0, 0, 0};
    SQ_CUDA(cudaMemcpyAsync(h_stats, c->w_stats.p, sizeof(h_stats), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[8], c->ev[9]);
    c->tm.walk_ms = ms;
    c->walk_done = true;
    c->walk_ntargets = ntargets;
    if(stats) {
        stats->ntargets = ntargets;
        stats->ngroups = (ntargets + G - 1) / G;
        stats->interactions = (int64_t) h_stats[0];
        stats->list_entries = (int64_t) h_stats[1];
        stats->nodes_tested = (int64_t) h_stats[2];
        stats->kernel_ms = ms;
    }
    // TreeUseBH > 1: Barnes-Hut on the first call only (gravshort-tree2.cpp:170-171)
    if(c->par.TreeUseBH > 1) c->par.TreeUseBH = 0;
    return SQ_OK;
}
