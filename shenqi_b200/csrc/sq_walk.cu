// Short-range tree walk: the interaction kernel.
//
// Replaces GravLocalTreeWalk::visit<PRIMARY> / apply_accn / shall_we_discard_node / shall_we_open_node
// (libgadget/gravshort2.hpp:152-358), TreeWalk::ev_primary (treewalk2.h:566-595), the GPU kernels
// treewalk_primary_kernel / treewalk_postprocess_kernel (treewalk2.cuh:104-173) and
// GravTreeOutput::postprocess (gravshort2.hpp:88-107).
//
// Design (not the reference's thread-per-particle depth-first walk):
//   * One warp walks the tree for a GROUP of 32 Peano-Hilbert-adjacent active particles with ONE node stack in
//     shared memory.  A stack entry is (node, mask): mask = the targets of the group that reach this node, i.e.
//     opened every ancestor.  Per step the warp pops up to 32 entries, one node per lane (three 128-bit loads).
//   * Every (target, node) decision is the reference's own (discard: the Rcut + len/2 box; open: relative
//     criterion M l^2 > r^4 alpha |a_old|, Barnes-Hut angle, inside-0.6-len box), reached in tiers so that FP64 is
//     the exception:
//       1. group tier (FP32 interval arithmetic over the group's bounding box and its min/max |a_old|);
//       1b. keep/discard for boundary nodes by BUCKET MASKS: per axis the targets are binned into 64 buckets over
//          the group's extent and S[i] = mask of the targets in buckets below i is tabulated once per group;
//          the targets inside [c - eff, c + eff] are then two table reads per bound, exact up to the targets in
//          the two boundary buckets, which fall to tier 2;
//       2. target tier (FP32, coordinates relative to the group centre in units of the split-table step, masses in
//          units of the mean particle mass): the class is evaluated at both ends of the rounding-error interval
//          of (r2, max-axis distance); equal -> decided;
//       3. exact tier: the reference's FP64 expressions with per-target NEAREST() for the pairs whose interval
//          straddles a threshold, and for nodes half a box away from the group.
//     The decision function is monotone in (r2, emax), so equal classes at the interval ends prove the FP64
//     class.  The interaction count therefore equals the reference's exactly.
//   * Accepted nodes and the particles of opened leaves go to FOUR interaction lists, one per OCTET of targets
//     (lanes 8q..8q+7), each with an 8-bit mask: an octet's bounding box is much smaller than the group's, so
//     the masked evaluation wastes fewer lanes (62 % useful instead of 41 %) while the walk is still shared by 32
//     targets.  The four octets evaluate their own lists in lock step: one 128-bit LDS with one address per
//     quarter warp costs the same as a warp-wide broadcast (measured, tools/ubench).
//   * Evaluation is FP32: rsqrt, the erfc/"exact" split factor from a shared-memory table REPLICATED 16x so that
//     every lane of a half warp reads its own bank pair (conflict-free: 2.2 clk per lookup instead of 4.9,
//     measured).  Distances are already in table units, so the table coordinate costs nothing.  Sources that
//     may come within the softening length of a target (inside the group's box grown by h), and everything when
//     the periodic image can differ per pair (tiny boxes), go to a short "near" list evaluated by a careful
//     loop (spline softening gravshort2.hpp:332-350, self pair r = 0, per-pair wrap); the fast loop has neither.
//     Partial sums are FP32 for one list flush (<= 64 entries), then added to FP64 accumulators.
//   * Epilogue = postprocess: x G, potential self terms; results written in sorted order.
#include "sq_common.cuh"
#include "shortrange_table.h"
#include <math.h>

#define WALK_WARPS 16
#define WALK_THREADS (WALK_WARPS * 32)
#define STACK_CAP 384
#define STACK_RESERVE 160   /* 7 siblings x 22 levels: room a depth-first descent can always need */
#define OL_CAP 64           /* entries per octet list (multiple of 4, >= 32) */
#define NEAR_CAP 32
#define NBUCK 64            /* buckets per axis over the group's extent */
#define NCUM (NBUCK + 5)
#define TAB_ROWS (SQ_NTAB + 1)
#define TAB_COPIES 16
#define TAB_BYTES (TAB_ROWS * TAB_COPIES * 8)
#define SQ_EPS 4e-6f        /* relative slack on every FP32 threshold (>> 2^-24 rounding chains) */
#define SQ_FAR 3.0e4f       /* padding entries: far beyond the table, zero mask */
#define SQ_MAGIC 12582912.f /* 1.5 * 2^23: adding it rounds to the nearest integer, kept in the mantissa */

struct sq_walk_par {
    // exact tier (reference expressions, FP64)
    double box, half, rcut, rcut2, bh2, h, errtol;
    // scaling: lengths x s (table units), masses x minv
    double s, minv, Ka, Kp;       // accel = Ka * sum, potential = Kp * sum (+ self terms)
    double G, cbrtrho0;
    int relative;                 // TreeUseBH == 0
    int gw;                       // targets per warp: 32, 16 or 8 (lanes beyond stay idle)
    // FP32 tiers (table units)
    float f_rcut, f_rc2_lo, f_rc2_hi, f_bh2, f_h2, f_hs, f_boxs, f_invboxs, f_halfs, f_range;
};

struct sq_walk_warp_smem {
    float4 ol[4][OL_CAP];         // octet interaction lists: x, y, z relative to the group centre (table units), mass
    float4 nl[NEAR_CAP];          // near list (careful loop)
    float4 tgt[32];               // targets: x, y, z, errtol*|a_old| scaled
    int2 stack[STACK_CAP];
    unsigned nm[NEAR_CAP];
    unsigned cum[3][NCUM + 3];    // bucket prefix masks S[i] = targets with bucket < i - 1
    unsigned char om[4][OL_CAP];  // 8-bit masks of the octet lists
    int s_end[32], s_first[32];
    unsigned s_mask[32];
};

__device__ __forceinline__ float2 sq_lds_f2(unsigned addr)
{
    float2 v;       // only used on the split table, constant after the kernel prologue
    asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}
// one MUFU.RSQ (the caller keeps the argument away from 0 and denormals)
__device__ __forceinline__ float sq_rsqrt(float x)
{
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// spline-softened force and potential factors, gravshort2.hpp:332-350 (FP64: rare branch).
// r, h in table units; returns (fac, facpot) per unit mass in the scaled system.
__device__ __noinline__ float2 sq_soft_kernel(float r, float hf)
{
    const double h = (double) hf;
    const double h_inv = 1.0 / h;
    const double h3_inv = h_inv * h_inv * h_inv;
    const double u = (double) r * h_inv;
    double f, wp;
    if(u < 0.5) {
        f = h3_inv * (10.666666666667 + u * u * (32.0 * u - 38.4));
        wp = -2.8 + u * u * (5.333333333333 + u * u * (6.4 * u - 9.6));
    } else {
        f = h3_inv * (21.333333333333 - 48.0 * u + 38.4 * u * u - 10.666666666667 * u * u * u - 0.066666666667 / (u * u * u));
        wp = -3.2 + 0.066666666667 / u + u * u * (10.666666666667 + u * (-16.0 + u * (9.6 - 2.133333333333 * u)));
    }
    return make_float2((float) f, (float) (h_inv * wp));
}

// exact tier: the reference's FP64 test for one (target, node)
struct sq_exact_par { double box, half, rcut, rcut2, bh2; int relative; };
__device__ __noinline__ int sq_exact_class(const double4 M, const double4 C, double x, double y, double z, double aold, const sq_exact_par P)
{
    const double len = C.w;
    const double dx = sq_nearest(M.x - x, P.box, P.half);
    const double dy = sq_nearest(M.y - y, P.box, P.half);
    const double dz = sq_nearest(M.z - z, P.box, P.half);
    const double r2 = dx * dx + dy * dy + dz * dz;
    const double ex = fabs(sq_nearest(C.x - x, P.box, P.half));
    const double ey = fabs(sq_nearest(C.y - y, P.box, P.half));
    const double ez = fabs(sq_nearest(C.z - z, P.box, P.half));
    // shall_we_discard_node, gravshort2.hpp:152-167
    if(r2 > P.rcut2) {
        const double eff = P.rcut + 0.5 * len;
        if(ex > eff || ey > eff || ez > eff) return 0;
    }
    // shall_we_open_node, gravshort2.hpp:172-193
    if(P.relative && (M.w * len * len > r2 * r2 * aold)) return 2;
    if(len * len / r2 > P.bh2) return 2;
    const double inside = 0.6 * len;
    if(ex < inside && ey < inside && ez < inside) return 2;
    return 1;
}

// bucket of coordinate x (monotone non-decreasing in x): 0 below the group's extent, 1..NBUCK inside, NBUCK+1 above
__device__ __forceinline__ int sq_bucket(float x, float inv, float off)
{
    return __float2int_rz(fminf(fmaxf(fmaf(x, inv, off), 0.f), (float) (NBUCK + 1)));
}

// ---- fast evaluation: the four octets run their own lists in lock step (lane = target 8q + lane%8).
// Every genuine entry is farther than the softening length from every target of the group (the others are in
// the near list), padding entries are SQ_FAR away, so r2 > 0 and no softening / wrap test is needed here.
// Table: (gravity.h:50-60) i = floor(r), linear interpolation, zero from row 511 on.
//   !POT: row i = (value at the middle of [i, i+1], slope); row 511 = (0, 0)
//    POT: row i = (force[i], potential[i]); rows i and i+1 are read; the r < 511 test zeroes the tail
template <bool POT>
__device__ __forceinline__ void sq_eval_octets(const sq_walk_warp_smem & W, int nmax4, unsigned tabaddr, float tx, float ty, float tz,
                                               int q, unsigned bit0, float & ax, float & ay, float & az, float & ap)
{
    const float4 * lst = W.ol[q];
    const unsigned * mk = reinterpret_cast<const unsigned *>(W.om[q]);
    const int nb = nmax4 >> 2;
    for(int b = 0; b < nb; b++) {
        const unsigned m4 = mk[b];
        #pragma unroll
        for(int j = 0; j < 4; j++) {
            const float4 s = lst[4 * b + j];
            const float dx = s.x - tx, dy = s.y - ty, dz = s.z - tz;
            const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            const bool on = (m4 & (bit0 << (8 * j))) != 0u;
            const float rinv = sq_rsqrt(r2);
            // xq = r - 1/2; round-to-nearest of xq = floor(r); d = r - floor(r) - 1/2
            const float xq = fminf(fmaf(r2, rinv, -0.5f), 511.f);
            const float t = xq + SQ_MAGIC;
            const float d = xq - (t - SQ_MAGIC);
            const unsigned addr = ((unsigned) __float_as_int(t) << 7) + tabaddr;
            const float rinv2 = rinv * rinv;
            if(POT) {
                const float2 e0 = sq_lds_f2(addr), e1 = sq_lds_f2(addr + 128);
                const float fr = d + 0.5f;
                const float wf = fmaf(fr, e1.x - e0.x, e0.x), wp = fmaf(fr, e1.y - e0.y, e0.y);
                const float mm = (on && r2 < 261121.f) ? s.w : 0.f;
                const float fac = (mm * rinv2) * (rinv * wf);
                ax = fmaf(dx, fac, ax);
                ay = fmaf(dy, fac, ay);
                az = fmaf(dz, fac, az);
                ap = fmaf(-mm * rinv, wp, ap);
            } else {
                const float2 w = sq_lds_f2(addr);
                const float mm = on ? s.w : 0.f;
                const float fac = (mm * rinv2) * (rinv * fmaf(d, w.y, w.x));
                ax = fmaf(dx, fac, ax);
                ay = fmaf(dy, fac, ay);
                az = fmaf(dz, fac, az);
            }
        }
    }
}

// ---- careful evaluation of the near list (lane = target, entries broadcast): softened pairs (r < 2.8 eps),
// the target meeting itself in its own leaf (r = 0, as in the reference), per-pair periodic wrap.
template <bool POT>
__device__ __forceinline__ void sq_eval_near(const sq_walk_warp_smem & W, int nn, unsigned tabrow0, const sq_walk_par & P, bool pairwrap,
                                             float tx, float ty, float tz, unsigned lane, float & ax, float & ay, float & az, float & ap)
{
    for(int k = 0; k < nn; k++) {
        const float4 s = W.nl[k];
        const unsigned m = W.nm[k];
        if((m >> lane) & 1u) {
            float dx = s.x - tx, dy = s.y - ty, dz = s.z - tz;
            if(pairwrap) {
                dx -= P.f_boxs * rintf(dx * P.f_invboxs);
                dy -= P.f_boxs * rintf(dy * P.f_invboxs);
                dz -= P.f_boxs * rintf(dz * P.f_invboxs);
            }
            const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            float r, fac, facpot;
            if(r2 >= P.f_h2) {
                const float rinv = sq_rsqrt(r2);
                r = r2 * rinv; fac = rinv * rinv * rinv; facpot = -rinv;
            } else {
                r = sqrtf(r2);
                const float2 sf = sq_soft_kernel(r, P.f_hs);
                fac = sf.x; facpot = sf.y;
            }
            if(r < 511.f) {
                const float fl = floorf(r);
                const float fr = r - fl;
                const unsigned addr = tabrow0 + ((unsigned) (int) fl << 7);
                float wf, wp = 0.f;
                if(POT) {
                    const float2 e0 = sq_lds_f2(addr), e1 = sq_lds_f2(addr + 128);
                    wf = fmaf(fr, e1.x - e0.x, e0.x); wp = fmaf(fr, e1.y - e0.y, e0.y);
                } else {
                    const float2 w = sq_lds_f2(addr);
                    wf = fmaf(fr - 0.5f, w.y, w.x);
                }
                const float f = s.w * fac * wf;
                ax = fmaf(dx, f, ax);
                ay = fmaf(dy, f, ay);
                az = fmaf(dz, f, az);
                if(POT) ap = fmaf(s.w * facpot, wp, ap);
            }
        }
    }
}

template <bool POT>
__global__ void __launch_bounds__(WALK_THREADS, 1)
sq_walk_kernel(const double4 * __restrict__ cm, const double4 * __restrict__ cl, const int4 * __restrict__ lk,
               const double * __restrict__ spos, const float * __restrict__ smass,
               const double * __restrict__ tpos, const float * __restrict__ tmass, const double * __restrict__ toldacc,
               const uint32_t * __restrict__ targets, long long ntargets, const float2 * __restrict__ gtab,
               double * __restrict__ accel, double * __restrict__ pot, unsigned long long * __restrict__ stats,
               int * __restrict__ group_counter, const sq_walk_par P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2 * tab = reinterpret_cast<float2 *>(smem_raw);
    sq_walk_warp_smem * Wall = reinterpret_cast<sq_walk_warp_smem *>(smem_raw + TAB_BYTES);
    // 16 copies of every table row side by side: copy c of row i at byte i*128 + c*8
    for(int i = threadIdx.x; i < TAB_ROWS * TAB_COPIES; i += blockDim.x)
        tab[i] = gtab[i / TAB_COPIES];
    const unsigned lane = sq_lane();
    sq_walk_warp_smem & W = Wall[threadIdx.x >> 5];
    __syncthreads();
    const unsigned tabrow0 = (unsigned) __cvta_generic_to_shared(tab) + (lane & 15u) * 8u;
    // biased so that the mantissa bits of (x + 1.5*2^23) index the rows directly
    const unsigned tabaddr = tabrow0 - (0x4B400000u << 7);
    const int q = (int) (lane >> 3);
    const unsigned bit0 = 1u << (lane & 7u);

    const int G = P.gw;
    const long long ngroups = (ntargets + G - 1) / G;
    unsigned long long st_int = 0, st_ent = 0, st_tst = 0, st_pair = 0, st_exact = 0;

    for(;;) {
        long long g = 0;
        if(lane == 0) g = atomicAdd(group_counter, 1);
        g = __shfl_sync(SQ_FULL, g, 0);
        if(g >= ngroups) break;
        const long long t0 = g * G;
        const int nt = (int) ((ntargets - t0) < G ? (ntargets - t0) : G);
        const bool valid = (int) lane < nt;
        long long tidx = 0;
        double px = 0, py = 0, pz = 0, aold = 0;
        if(valid) {
            tidx = targets ? (long long) targets[t0 + lane] : (t0 + lane);
            px = tpos[3 * tidx]; py = tpos[3 * tidx + 1]; pz = tpos[3 * tidx + 2];
            aold = P.errtol * toldacc[tidx];
        }
        // group bounding box over the valid targets
        double lox = valid ? px : 1e300, loy = valid ? py : 1e300, loz = valid ? pz : 1e300;
        double hix = valid ? px : -1e300, hiy = valid ? py : -1e300, hiz = valid ? pz : -1e300;
        double amin = valid ? aold : 1e300, amax = valid ? aold : 0.0;
        #pragma unroll
        for(int o = 16; o >= 1; o >>= 1) {
            lox = fmin(lox, __shfl_xor_sync(SQ_FULL, lox, o)); hix = fmax(hix, __shfl_xor_sync(SQ_FULL, hix, o));
            loy = fmin(loy, __shfl_xor_sync(SQ_FULL, loy, o)); hiy = fmax(hiy, __shfl_xor_sync(SQ_FULL, hiy, o));
            loz = fmin(loz, __shfl_xor_sync(SQ_FULL, loz, o)); hiz = fmax(hiz, __shfl_xor_sync(SQ_FULL, hiz, o));
            amin = fmin(amin, __shfl_xor_sync(SQ_FULL, amin, o)); amax = fmax(amax, __shfl_xor_sync(SQ_FULL, amax, o));
        }
        const double gcx = 0.5 * (lox + hix), gcy = 0.5 * (loy + hiy), gcz = 0.5 * (loz + hiz);
        // half extents in table units, rounded up
        const float ghx = (float) (0.5 * (hix - lox) * P.s) * 1.000001f + 1e-30f;
        const float ghy = (float) (0.5 * (hiy - loy) * P.s) * 1.000001f + 1e-30f;
        const float ghz = (float) (0.5 * (hiz - loz) * P.s) * 1.000001f + 1e-30f;
        const float ghm = fmaxf(ghx, fmaxf(ghy, ghz));
        // errtol*|a_old| in the scaled system: M l^2 > r^4 a  <=>  M' l'^2 > r'^4 a / (munit s^2); BH mode: never
        const double ascale = P.minv / (P.s * P.s);
        const float ta = P.relative ? (float) (aold * ascale) : __int_as_float(0x7f800000);
        const float ta_lo = P.relative ? (float) (amin * ascale) * (1.f - SQ_EPS) : __int_as_float(0x7f800000);
        const float ta_hi = P.relative ? (float) (amax * ascale) * (1.f + SQ_EPS) : __int_as_float(0x7f800000);
        // idle lanes sit far away so that they never see r = 0
        const float tx = valid ? (float) ((px - gcx) * P.s) : 2.f * SQ_FAR;
        const float ty = valid ? (float) ((py - gcy) * P.s) : 2.f * SQ_FAR;
        const float tz = valid ? (float) ((pz - gcz) * P.s) : 2.f * SQ_FAR;
        __syncwarp();
        W.tgt[lane] = make_float4(tx, ty, tz, ta);
        // ---- bucket prefix masks, per axis: S[i] = targets whose bucket is below i - 1
        const float bwx = fmaxf(ghx, 1e-3f), bwy = fmaxf(ghy, 1e-3f), bwz = fmaxf(ghz, 1e-3f);
        const float invx = (0.999f * NBUCK) / (2.f * bwx), invy = (0.999f * NBUCK) / (2.f * bwy), invz = (0.999f * NBUCK) / (2.f * bwz);
        const float offx = fmaf(bwx, invx, 1.f), offy = fmaf(bwy, invy, 1.f), offz = fmaf(bwz, invz, 1.f);
        // a bound is resolved to +-1 bucket only while the FP32 uncertainty stays below 0.9 bucket widths
        const float wmin = 0.9f / fmaxf(invx, fmaxf(invy, invz));
        for(int i = lane; i < 3 * (NCUM + 3); i += 32)
            (&W.cum[0][0])[i] = 0u;
        __syncwarp();
        if(valid) {
            atomicOr(&W.cum[0][sq_bucket(tx, invx, offx) + 2], 1u << lane);
            atomicOr(&W.cum[1][sq_bucket(ty, invy, offy) + 2], 1u << lane);
            atomicOr(&W.cum[2][sq_bucket(tz, invz, offz) + 2], 1u << lane);
        }
        __syncwarp();
        #pragma unroll
        for(int a = 0; a < 3; a++) {
            const int i0 = 3 * (int) lane;
            unsigned v0 = 0, v1 = 0, v2 = 0;
            if(i0 + 2 < NCUM + 3) { v0 = W.cum[a][i0]; v1 = v0 | W.cum[a][i0 + 1]; v2 = v1 | W.cum[a][i0 + 2]; }
            unsigned x = v2;
            #pragma unroll
            for(int o = 1; o < 32; o <<= 1) {
                const unsigned y = __shfl_up_sync(SQ_FULL, x, o);
                if((int) lane >= o) x |= y;
            }
            unsigned ex = __shfl_up_sync(SQ_FULL, x, 1);
            if(lane == 0) ex = 0u;
            if(i0 + 2 < NCUM + 3) { W.cum[a][i0] = ex | v0; W.cum[a][i0 + 1] = ex | v1; W.cum[a][i0 + 2] = ex | v2; }
        }
        // sources inside this box may come within the softening length of a target -> near list
        const float nhx = ghx + P.f_hs * 1.001f + 1e-3f, nhy = ghy + P.f_hs * 1.001f + 1e-3f, nhz = ghz + P.f_hs * 1.001f + 1e-3f;
        // pairs may straddle the periodic image once group extent + table range reaches half the box
        const bool pairwrap = 2.f * ghm + P.f_range >= P.f_halfs;
        float ax = 0.f, ay = 0.f, az = 0.f, ap = 0.f;
        double dax = 0.0, day = 0.0, daz = 0.0, dap = 0.0;
        int top = 1, n0 = 0, n1 = 0, n2 = 0, n3 = 0, nn = 0;
        if(lane == 0) W.stack[0] = make_int2(0, nt >= 32 ? (int) 0xffffffffu : (int) ((1u << nt) - 1u));
        __syncwarp();

        for(;;) {
            // the last pass (empty stack) only drains the lists
            const bool drain = top == 0;
            unsigned acc_mask = 0, open_mask = 0;
            int4 L = make_int4(0, 0, -1, 0);
            float4 nodeE = make_float4(0.f, 0.f, 0.f, 0.f);
            if(!drain) {
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
                if(have) {
                    L = lk[node];
                    const double4 M = cm[node];
                    const double4 C = cl[node];
                    st_tst++;
                    const double dmx = sq_nearest(M.x - gcx, P.box, P.half), dmy = sq_nearest(M.y - gcy, P.box, P.half), dmz = sq_nearest(M.z - gcz, P.box, P.half);
                    const double dcx = sq_nearest(C.x - gcx, P.box, P.half), dcy = sq_nearest(C.y - gcy, P.box, P.half), dcz = sq_nearest(C.z - gcz, P.box, P.half);
                    const float cmx = (float) (dmx * P.s), cmy = (float) (dmy * P.s), cmz = (float) (dmz * P.s);
                    const float cx = (float) (dcx * P.s), cy = (float) (dcy * P.s), cz = (float) (dcz * P.s);
                    const float n_mass = (float) (M.w * P.minv);
                    nodeE = make_float4(cmx, cmy, cmz, n_mass);
                    const float len = (float) (C.w * P.s);
                    const float l2 = len * len, ml2 = n_mass * l2;
                    const float acx = fabsf(cx), acy = fabsf(cy), acz = fabsf(cz);
                    const float amx = fabsf(cmx), amy = fabsf(cmy), amz = fabsf(cmz);
                    // rounding-error bounds of the FP32 coordinates / squared distances
                    const float scale = fmaxf(fmaxf(fmaxf(acx, acy), fmaxf(acz, amx)), fmaxf(amy, amz)) + ghm;
                    const float de = 3e-7f * scale;
                    const float R1 = amx + amy + amz + 3.f * ghm;
                    const float dr2 = 4.f * R1 * de + 1e-6f * R1 * R1;
                    const float eff = P.f_rcut + 0.5f * len, ins = 0.6f * len;
                    // thresholds with the relative slack and the coordinate error folded in:
                    //   "A" = the end of the interval that keeps / opens most, "B" = the end that keeps / opens least
                    const float rcA = P.f_rc2_hi + dr2, rcB = P.f_rc2_lo - dr2;
                    const float efA = eff * (1.f + SQ_EPS) + de, efB = eff * (1.f - SQ_EPS) - de;
                    const float inA = ins * (1.f + SQ_EPS) + de, inB = ins * (1.f - SQ_EPS) - de;
                    const float bhq = l2 / P.f_bh2;
                    const float bhA = bhq * (1.f + SQ_EPS) + dr2, bhB = bhq * (1.f - SQ_EPS) - dr2;
                    const float mlA = ml2 * (1.f + SQ_EPS), mlB = ml2 * (1.f - SQ_EPS);
                    // a node about half a box away from the group: per-target NEAREST may differ from the group's
                    const bool risky = fmaxf(fmaxf(acx + ghx, acy + ghy), fmaxf(acz + ghz, fmaxf(fmaxf(amx + ghx, amy + ghy), amz + ghz))) >= P.f_halfs * 0.999f;
                    unsigned exact = 0, kept = 0, amb = 0;
                    if(risky) exact = mask;
                    else {
                        // ---- keep or discard (shall_we_discard_node).  With the centre of mass inside the cell,
                        // "r2 <= Rcut^2" implies the box test, so the box decides; otherwise every target is tested.
                        const float en = fmaxf(fmaxf(acx - ghx, acy - ghy), acz - ghz);
                        const float ef = fmaxf(fmaxf(acx + ghx, acy + ghy), acz + ghz);
                        const float hl = 0.5f * len * (1.f + 1e-5f) + de;
                        const bool cmin = fabsf(cmx - cx) <= hl && fabsf(cmy - cy) <= hl && fabsf(cmz - cz) <= hl;
                        if(!cmin) amb = mask;
                        else if(en > efA) { }                       // outside every target's box
                        else if(!(ef > efB)) kept = mask;           // inside every target's box
                        else if(de + eff * SQ_EPS < wmin) {
                            // tier 1b: bucket masks, exact up to the targets in the boundary buckets
                            const int bxl = sq_bucket(cx - eff, invx, offx), bxu = sq_bucket(cx + eff, invx, offx);
                            const int byl = sq_bucket(cy - eff, invy, offy), byu = sq_bucket(cy + eff, invy, offy);
                            const int bzl = sq_bucket(cz - eff, invz, offz), bzu = sq_bucket(cz + eff, invz, offz);
                            const unsigned sure = ~W.cum[0][bxl + 3] & W.cum[0][bxu] & ~W.cum[1][byl + 3] & W.cum[1][byu] & ~W.cum[2][bzl + 3] & W.cum[2][bzu];
                            const unsigned maybe = ~W.cum[0][bxl] & W.cum[0][bxu + 3] & ~W.cum[1][byl] & W.cum[1][byu + 3] & ~W.cum[2][bzl] & W.cum[2][bzu + 3];
                            kept = sure & mask;
                            amb = maybe & ~sure & mask;
                        } else amb = mask;
                        // ---- tier 2a: keep or discard, per target, FP32 with error intervals
                        st_pair += __popc(amb);
                        while(amb) {
                            const int t = 31 - __clz(amb);
                            const unsigned bit = 1u << t;
                            amb ^= bit;
                            const float4 T = W.tgt[t];
                            const float dx = cmx - T.x, dy = cmy - T.y, dz = cmz - T.z;
                            const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                            const float em = fmaxf(fmaxf(fabsf(cx - T.x), fabsf(cy - T.y)), fabsf(cz - T.z));
                            const bool k_lo = (r2 <= rcA) || !(em > efA);
                            const bool k_hi = (r2 <= rcB) || !(em > efB);
                            if(k_lo != k_hi) exact |= bit;
                            else if(k_lo) kept |= bit;
                        }
                        // ---- open or accept (shall_we_open_node): the whole group at once if both ends of the
                        // interval over the group's box and min/max |a_old| agree, else tier 2b per target
                        if(kept) {
                            const float qx = fmaxf(amx - ghx, 0.f), qy = fmaxf(amy - ghy, 0.f), qz = fmaxf(amz - ghz, 0.f);
                            const float wx = amx + ghx, wy = amy + ghy, wz = amz + ghz;
                            const float r2n = qx * qx + qy * qy + qz * qz, r2f = wx * wx + wy * wy + wz * wz;
                            const float r2nl = fmaxf(r2n - dr2, 0.f), r2fh = r2f + dr2;
                            const bool go_lo = (mlA > r2nl * r2nl * ta_lo) || (r2n < bhA) || (en < inA);
                            const bool go_hi = (mlB > r2fh * r2fh * ta_hi) || (r2f < bhB) || (ef < inB);
                            if(go_lo == go_hi) {
                                if(go_hi) open_mask = kept; else acc_mask = kept;
                            } else {
                                st_pair += __popc(kept);
                                while(kept) {
                                    const int t = 31 - __clz(kept);
                                    const unsigned bit = 1u << t;
                                    kept ^= bit;
                                    const float4 T = W.tgt[t];
                                    const float dx = cmx - T.x, dy = cmy - T.y, dz = cmz - T.z;
                                    const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                                    const float em = fmaxf(fmaxf(fabsf(cx - T.x), fabsf(cy - T.y)), fabsf(cz - T.z));
                                    const float r2l = fmaxf(r2 - dr2, 0.f), r2h = r2 + dr2;
                                    const bool o_lo = (mlA > r2l * r2l * T.w) || (r2 < bhA) || (em < inA);
                                    const bool o_hi = (mlB > r2h * r2h * T.w) || (r2 < bhB) || (em < inB);
                                    if(o_lo != o_hi) exact |= bit;
                                    else if(o_lo) open_mask |= bit;
                                    else acc_mask |= bit;
                                }
                            }
                        }
                    }
                    // ---- tier 3: the reference's FP64 expressions
                    while(exact) {
                        const int t = 31 - __clz(exact);
                        const unsigned bit = 1u << t;
                        exact ^= bit;
                        const long long ti = targets ? (long long) targets[t0 + t] : (t0 + t);
                        sq_exact_par X;
                        X.box = P.box; X.half = P.half; X.rcut = P.rcut; X.rcut2 = P.rcut2; X.bh2 = P.bh2; X.relative = P.relative;
                        const int c = sq_exact_class(M, C, tpos[3 * ti], tpos[3 * ti + 1], tpos[3 * ti + 2], P.errtol * toldacc[ti], X);
                        st_exact++;
                        if(c == 1) acc_mask |= bit;
                        else if(c == 2) open_mask |= bit;
                    }
                }
            }
            const int type = SQ_INFO_TYPE(L.w);
            // ---- opened internal nodes -> push children with the open mask (pseudo nodes are skipped: gravshort2.hpp:305-314)
            {
                const int nch = (open_mask && type == 1) ? L.y : 0;
                int tot;
                const int off = sq_warp_excl_scan(nch, tot);
                for(int j = 0; j < nch; j++)
                    W.stack[top + off + j] = make_int2(L.x + j, (int) open_mask);
                top += tot;
            }
            // ---- opened leaves -> their particles are sources (gravshort2.hpp:291-303)
            int tot_leaf;
            {
                const int np = (open_mask && type == 0) ? L.y : 0;
                const int off = sq_warp_excl_scan(np, tot_leaf);
                if(tot_leaf) {
                    st_int += (unsigned long long) np * __popc(open_mask);
                    W.s_end[lane] = off + np;
                    W.s_first[lane] = L.x - off;        // particle = s_first + j
                    W.s_mask[lane] = open_mask;
                }
            }
            st_int += __popc(acc_mask);
            const unsigned anyacc = __ballot_sync(SQ_FULL, acc_mask != 0u);
            __syncwarp();
            // ---- sources join the lists, 32 per round: round 0 = accepted nodes, then the leaf particles
            const int nrounds = 1 + ((tot_leaf + 31) >> 5);
            for(int rd = 0; rd < nrounds; rd++) {
                bool v = false;
                float4 e = nodeE;
                unsigned m = acc_mask;
                if(rd == 0) {
                    if(!drain && anyacc == 0u) continue;
                    v = acc_mask != 0u;
                } else {
                    const int j = ((rd - 1) << 5) + (int) lane;
                    if(j < tot_leaf) {
                        // first source lane whose end is past j
                        int lo = 0, hi = 31;
                        while(lo < hi) {
                            const int mid = (lo + hi) >> 1;
                            if(W.s_end[mid] > j) hi = mid; else lo = mid + 1;
                        }
                        const long long p = (long long) W.s_first[lo] + j;
                        const double sx = spos[3 * p], sy = spos[3 * p + 1], sz = spos[3 * p + 2];
                        e = make_float4((float) (sq_nearest(sx - gcx, P.box, P.half) * P.s), (float) (sq_nearest(sy - gcy, P.box, P.half) * P.s),
                                        (float) (sq_nearest(sz - gcz, P.box, P.half) * P.s), (float) ((double) smass[p] * P.minv));
                        m = W.s_mask[lo];
                        v = true;
                    }
                }
                // ---- near list
                const bool near = v && (pairwrap || (fabsf(e.x) <= nhx && fabsf(e.y) <= nhy && fabsf(e.z) <= nhz));
                const unsigned bn = __ballot_sync(SQ_FULL, near);
                if(bn != 0u || (drain && nn > 0)) {
                    const int cn = __popc(bn);
                    if(nn + cn > NEAR_CAP || drain) {
                        __syncwarp();
                        sq_eval_near<POT>(W, nn, tabrow0, P, pairwrap, tx, ty, tz, lane, ax, ay, az, ap);
                        st_ent += nn;
                        nn = 0;
                        __syncwarp();
                    }
                    if(near) {
                        const int p = nn + __popc(bn & sq_lanemask_lt());
                        W.nl[p] = e;
                        W.nm[p] = m;
                    }
                    nn += cn;
                }
                // ---- octet lists
                const bool isfar = v && !near;
                const unsigned b0 = __ballot_sync(SQ_FULL, isfar && (m & 0x000000ffu));
                const unsigned b1 = __ballot_sync(SQ_FULL, isfar && (m & 0x0000ff00u));
                const unsigned b2 = __ballot_sync(SQ_FULL, isfar && (m & 0x00ff0000u));
                const unsigned b3 = __ballot_sync(SQ_FULL, isfar && (m & 0xff000000u));
                const int c0 = __popc(b0), c1 = __popc(b1), c2 = __popc(b2), c3 = __popc(b3);
                if(n0 + c0 > OL_CAP || n1 + c1 > OL_CAP || n2 + c2 > OL_CAP || n3 + c3 > OL_CAP || (drain && (n0 | n1 | n2 | n3) != 0)) {
                    const int nmax = max(max(n0, n1), max(n2, n3));
                    const int nmax4 = (nmax + 3) & ~3;
                    // pad the shorter lists up to the common length
                    const int nq = q == 0 ? n0 : (q == 1 ? n1 : (q == 2 ? n2 : n3));
                    for(int kk = nq + (int) (lane & 7u); kk < nmax4; kk += 8) {
                        W.ol[q][kk] = make_float4(SQ_FAR, SQ_FAR, SQ_FAR, 0.f);
                        W.om[q][kk] = 0;
                    }
                    __syncwarp();
                    sq_eval_octets<POT>(W, nmax4, tabaddr, tx, ty, tz, q, bit0, ax, ay, az, ap);
                    st_ent += nmax4;
                    dax += (double) ax; day += (double) ay; daz += (double) az;
                    ax = ay = az = 0.f;
                    if(POT) { dap += (double) ap; ap = 0.f; }
                    n0 = n1 = n2 = n3 = 0;
                    __syncwarp();
                }
                if(isfar) {
                    const unsigned lt = sq_lanemask_lt();
                    if(m & 0x000000ffu) { const int p = n0 + __popc(b0 & lt); W.ol[0][p] = e; W.om[0][p] = (unsigned char) (m & 0xffu); }
                    if(m & 0x0000ff00u) { const int p = n1 + __popc(b1 & lt); W.ol[1][p] = e; W.om[1][p] = (unsigned char) ((m >> 8) & 0xffu); }
                    if(m & 0x00ff0000u) { const int p = n2 + __popc(b2 & lt); W.ol[2][p] = e; W.om[2][p] = (unsigned char) ((m >> 16) & 0xffu); }
                    if(m & 0xff000000u) { const int p = n3 + __popc(b3 & lt); W.ol[3][p] = e; W.om[3][p] = (unsigned char) (m >> 24); }
                }
                n0 += c0; n1 += c1; n2 += c2; n3 += c3;
            }
            __syncwarp();
            if(drain) break;
        }
        dax += (double) ax; day += (double) ay; daz += (double) az;
        if(POT) dap += (double) ap;
        if(valid) {
            // gravshort2.hpp:88-107
            accel[3 * tidx] = dax * P.Ka;
            accel[3 * tidx + 1] = day * P.Ka;
            accel[3 * tidx + 2] = daz * P.Ka;
            if(POT) {
                const double mt = (double) tmass[tidx];
                double p = dap * P.Kp;
                p += mt / (P.h / 2.8);
                p -= 2.8372975 * pow(mt, 2.0 / 3) * P.cbrtrho0;
                pot[tidx] = p * P.G;
            }
        }
        __syncwarp();
    }
    // per-warp statistics
    #pragma unroll
    for(int o = 16; o >= 1; o >>= 1) {
        st_int += __shfl_xor_sync(SQ_FULL, st_int, o);
        st_tst += __shfl_xor_sync(SQ_FULL, st_tst, o);
        st_pair += __shfl_xor_sync(SQ_FULL, st_pair, o);
        st_exact += __shfl_xor_sync(SQ_FULL, st_exact, o);
    }
    if(lane == 0) {
        atomicAdd(&stats[0], st_int);
        atomicAdd(&stats[1], st_ent);
        atomicAdd(&stats[2], st_tst);
        atomicAdd(&stats[3], st_pair);
        atomicAdd(&stats[4], st_exact);
    }
}
// ------------------------------------------------------------------------------------ host side
static const double h_sr_x[SQ_NGRAVTAB] = SQ_SHORTRANGE_X_INIT;
static const double h_sr_pot[SQ_NGRAVTAB] = SQ_SHORTRANGE_POT_INIT;
static const double h_sr_force[SQ_NGRAVTAB] = SQ_SHORTRANGE_FORCE_INIT;

// GravShortTable, libgadget/gravshort-tree2.cpp:29-69: float tables; lookup gravity.h:50-60 (i = floor(r / dx),
// linear interpolation between rows i and i+1, nothing from row 511 on).  Two device forms, TAB_ROWS float2 rows:
//   force only:  row i = (force value at the middle of [i, i+1], slope), rows 511.. = (0, 0)
//   force + pot: row i = (force[i], potential[i]), row 512 = (0, 0)
static int sq_walk_table(sq_ctx * c)
{
    float ft[SQ_NTAB], pt[SQ_NTAB];
    for(int i = 0; i < SQ_NTAB; i++) {
        const double u = h_sr_x[i] * 0.5 / c->par.Asmth;
        if(c->have_raw) {
            ft[i] = c->raw_ft[i];
            pt[i] = c->raw_pt[i];
        } else if(c->par.WindowType == 1) {
            ft[i] = (float) h_sr_force[i];
            pt[i] = (float) h_sr_pot[i];
        } else {
            ft[i] = (float) (erfc(u) + 2.0 * u / sqrt(M_PI) * exp(-u * u));
            pt[i] = (float) erfc(u);
        }
    }
    float2 h[2][TAB_ROWS];
    for(int i = 0; i < TAB_ROWS; i++) {
        h[0][i] = make_float2(0.f, 0.f);
        h[1][i] = make_float2(0.f, 0.f);
        if(i + 1 < SQ_NTAB) {
            const double sf = (double) ft[i + 1] - (double) ft[i];
            h[0][i] = make_float2((float) ((double) ft[i] + 0.5 * sf), (float) sf);
        }
        if(i < SQ_NTAB) h[1][i] = make_float2(ft[i], pt[i]);
    }
    SQ_TRY(c->w_table.reserve(sizeof(h)));
    SQ_CUDA(cudaMemcpyAsync(c->w_table.p, h, sizeof(h), cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    return SQ_OK;
}

template <bool POT>
static int sq_walk_launch(sq_ctx * c, const uint32_t * targets, int64_t ntargets, const sq_walk_par & P)
{
    const size_t smem = TAB_BYTES + WALK_WARPS * sizeof(sq_walk_warp_smem);
    SQ_CUDA(cudaFuncSetAttribute(sq_walk_kernel<POT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    int per_sm = 0;
    SQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sq_walk_kernel<POT>, WALK_THREADS, smem));
    if(per_sm < 1) per_sm = 1;
    const int64_t ngroups = (ntargets + P.gw - 1) / P.gw;
    int64_t grid = (int64_t) c->num_sms * per_sm;
    const int64_t need = (ngroups + WALK_WARPS - 1) / WALK_WARPS;
    if(grid > need) grid = need;
    if(grid < 1) grid = 1;
    const double * spos = c->tree_identity ? c->pos.as<double>() : c->tree_pos.as<double>();
    const float * smass = c->tree_identity ? c->mass.as<float>() : c->tree_mass.as<float>();
    sq_walk_kernel<POT><<<(unsigned) grid, WALK_THREADS, smem, c->stream>>>(c->t_cm.as<double4>(), c->t_cl.as<double4>(), c->t_lk.as<int4>(),
            spos, smass, c->pos.as<double>(), c->mass.as<float>(), c->oldacc.as<double>(), targets, (long long) ntargets,
            c->w_table.as<float2>() + (POT ? TAB_ROWS : 0), c->w_accel.as<double>(), c->w_pot.as<double>(), c->w_stats.as<unsigned long long>(),
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
    double table_dx = SQ_SHORTRANGE_DX;
    double cell = cellsize;
    if(c->have_raw) {
        // the reference's own GravTreeParams (gravshort2.hpp:21-55), as handed to grav_short_tree_cuda
        const sq_walk_raw & R = c->raw;
        cell = R.cellsize; table_dx = R.table_dx;
        P.box = R.BoxSize; P.half = 0.5 * R.BoxSize;
        P.rcut = R.Rcut; P.rcut2 = R.Rcut * R.Rcut;
        P.h = R.ForceSoftening; P.G = R.G; P.cbrtrho0 = R.cbrtrho0; P.errtol = R.ErrTolForceAcc;
        P.relative = R.TreeUseBH == 0; P.bh2 = R.BHOpeningAngle2;
    }
    // scaled system of the FP32 tiers: 1 length unit = one step of the split table, 1 mass unit = mean tree particle
    P.s = 1.0 / (cell * table_dx);
    const double munit = (c->tree_total_mass > 0 && c->tree_np > 0) ? c->tree_total_mass / (double) c->tree_np : 1.0;
    P.minv = 1.0 / munit;
    P.Ka = P.G * munit * P.s * P.s;
    P.Kp = munit * P.s;
    P.gw = sq_walk_group_width;
    P.f_rcut = (float) (P.rcut * P.s);
    P.f_rc2_lo = (float) (P.rcut2 * P.s * P.s) * (1.f - SQ_EPS);
    P.f_rc2_hi = (float) (P.rcut2 * P.s * P.s) * (1.f + SQ_EPS);
    P.f_bh2 = (float) P.bh2;
    P.f_hs = (float) (P.h * P.s);
    P.f_h2 = P.f_hs * P.f_hs;
    P.f_boxs = (float) (P.box * P.s);
    P.f_invboxs = (float) (1.0 / (P.box * P.s));
    P.f_halfs = (float) (P.half * P.s);
    P.f_range = (float) (SQ_NTAB - 1);

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
    if(ntargets > 0)
        SQ_TRY(pot ? sq_walk_launch<true>(c, targets, ntargets, P) : sq_walk_launch<false>(c, targets, ntargets, P));
    SQ_CUDA(cudaEventRecord(c->ev[9], c->stream));
    unsigned long long h_stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    SQ_CUDA(cudaMemcpyAsync(h_stats, c->w_stats.p, sizeof(h_stats), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[8], c->ev[9]);
    c->tm.walk_ms = ms;
    c->walk_done = true;
    c->walk_ntargets = ntargets;
    if(stats) {
        stats->ntargets = ntargets;
        stats->ngroups = (ntargets + P.gw - 1) / P.gw;
        stats->interactions = (int64_t) h_stats[0];
        stats->list_entries = (int64_t) h_stats[1];
        stats->nodes_tested = (int64_t) h_stats[2];
        stats->pair_tests = (int64_t) h_stats[3];
        stats->exact_tests = (int64_t) h_stats[4];
        stats->kernel_ms = ms;
    }
    // TreeUseBH > 1: Barnes-Hut on the first call only (gravshort-tree2.cpp:170-171)
    if(c->par.TreeUseBH > 1 && !c->have_raw) c->par.TreeUseBH = 0;
    return SQ_OK;
}
