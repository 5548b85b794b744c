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
//   * Every (target, node) decision is the reference's own (discard: r2 vs Rcut and the Rcut+len/2 box;
//     open: relative criterion M l^2 > r^4 alpha |a_old|, Barnes-Hut angle, inside-0.6-len box), but it is
//     reached in three tiers so that FP64 is the exception:
//       1. group tier (FP32 interval arithmetic over the group's bounding box and its min/max |a_old|):
//          if both ends of the interval give the same class, the whole mask is discarded / accepted /
//          opened at once;
//       2. target tier (FP32, coordinates relative to the group centre in units of the split-table
//          spacing, masses in units of the mean particle mass): the class is evaluated at both ends of
//          the rounding-error interval of (r2, max-axis distance); equal -> decided;
//       3. exact tier: the reference's FP64 expressions with per-target NEAREST() for the pairs whose
//          interval straddles a threshold, and for nodes half a box away from the group.
//     The decision function is monotone in (r2, emax), so equal classes at the interval ends prove the
//     FP64 class.  The interaction count therefore equals the reference's exactly.
//   * Accepted nodes and the particles of opened leaves go to the interaction list with their masks;
//     children are pushed with the open mask.
//   * Evaluation is FP32: rsqrt, spline softening (rare, FP64), the erfc/"exact" split factor from a
//     shared-memory-resident table of (value at mid-point offset, slope) pairs -> one 64-bit LDS + one FMA.
//     Distances are already in table units, so the table coordinate costs nothing.  Partial sums are
//     carried in FP32 for one list flush (<= LIST_CAP entries) and then added to FP64 accumulators.
//   * Epilogue = postprocess: x G, potential self terms; results written in sorted order.
#include "sq_common.cuh"
#include "shortrange_table.h"
#include <math.h>

#define WALK_WARPS 8
#define WALK_THREADS (WALK_WARPS * 32)
#define STACK_CAP 512
#define STACK_RESERVE 160   /* 7 siblings x 22 levels: room a depth-first descent can always need */
#define LIST_CAP 128        /* multiple of 4 */
#define SQ_EPS 4e-6f        /* relative slack on every FP32 threshold (>> 2^-24 rounding chains) */

struct sq_walk_par {
    // exact tier (reference expressions, FP64)
    double box, half, rcut, rcut2, bh2, h, errtol;
    // scaling: lengths x s (table units), masses x minv
    double s, minv, Ka, Kp;       // accel = Ka * sum, potential = Kp * sum (+ self terms)
    double G, cbrtrho0;
    int relative;                 // TreeUseBH == 0
    int gw;                       // targets per warp: 32, 16 or 8
    // FP32 tiers (table units)
    float f_rcut, f_rc2_lo, f_rc2_hi, f_bh2, f_h2, f_hs, f_boxs, f_invboxs, f_halfs, f_range;
};

struct sq_walk_warp_smem {
    float4 tgt[32];               // x, y, z relative to the group centre (table units), w = errtol*|a_old| scaled
    int2 stack[STACK_CAP];
    float4 lst[LIST_CAP];
    unsigned lmask[LIST_CAP];
    int s_end[32], s_first[32];
    unsigned s_mask[32];
};

// spline-softened force and potential factors, gravshort2.hpp:332-350 (FP64: rare branch).
// r, h in table units; returns (fac, facpot) per unit mass in the scaled system.
__device__ __noinline__ float2 sq_soft_kernel(float r, double h)
{
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

template <bool POT>
struct sq_walk_acc {
    float ax, ay, az, ap;
    double dax, day, daz, dap;
};

__device__ __forceinline__ float2 sq_lds_f2(unsigned addr)
{
    float2 v;
    asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ float4 sq_lds_f4(unsigned addr)
{
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
// one MUFU.RSQ (the caller keeps the argument away from 0 and denormals)
__device__ __forceinline__ float sq_rsqrt(float x)
{
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Evaluate list entries [0, n) for this lane's target (slice sl of S takes the 4-entry blocks sl, sl+S, ...).
template <bool POT, bool WRAP>
__device__ __forceinline__ void sq_walk_flush(sq_walk_warp_smem & W, int n, unsigned tab_biased,
                                              const sq_walk_par & P, float tx, float ty, float tz, int tl, int sl, int S,
                                              sq_walk_acc<POT> & A)
{
    const int nb = (n + 3) >> 2;
    {
        const int pad = (nb << 2) - n;
        if((int) sq_lane() < pad) W.lmask[n + sq_lane()] = 0u;
    }
    __syncwarp();
    const unsigned lbit = 1u << tl;
    const uint4 * mk4 = reinterpret_cast<const uint4 *>(W.lmask);
    const float h2 = P.f_h2;
    for(int b = sl; b < nb; b += S) {
        const uint4 mk = mk4[b];
        unsigned soft = 0;
        #pragma unroll
        for(int j = 0; j < 4; j++) {
            const float4 s = W.lst[4 * b + j];
            const unsigned m = j == 0 ? mk.x : (j == 1 ? mk.y : (j == 2 ? mk.z : mk.w));
            float dx = s.x - tx, dy = s.y - ty, dz = s.z - tz;
            if(WRAP) {
                dx -= P.f_boxs * rintf(dx * P.f_invboxs);
                dy -= P.f_boxs * rintf(dy * P.f_invboxs);
                dz -= P.f_boxs * rintf(dz * P.f_invboxs);
            }
            const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            const bool on = (m & lbit) != 0u;
            const bool far = r2 >= h2;
            const float mm = (on && far) ? s.w : 0.f;
            if(on && !far) soft |= 1u << j;
            // softened pairs (r2 < h2) take the slow path below; clamping keeps rsqrt finite for them
            const float rinv = sq_rsqrt(fmaxf(r2, h2));
            // table coordinate: distances are in table units.  xq = r - 1/2; round-to-nearest of xq = floor(r)
            // (gravity.h:50-60: i = r/cellsize/dx, zero beyond the table, linear interpolation)
            const float xq = fminf(fmaf(r2, rinv, -0.5f), 511.f);
            const float t = xq + 12582912.f;
            const float d = xq - (t - 12582912.f);
            const unsigned off = (unsigned) __float_as_int(t) << 4;
            const float rinv2 = rinv * rinv;
            if(POT) {
                const float4 w = sq_lds_f4(tab_biased + off);
                const float fac = (mm * rinv2) * (rinv * fmaf(d, w.y, w.x));
                A.ax = fmaf(dx, fac, A.ax);
                A.ay = fmaf(dy, fac, A.ay);
                A.az = fmaf(dz, fac, A.az);
                A.ap = fmaf(-mm * rinv, fmaf(d, w.w, w.z), A.ap);
            } else {
                const float2 w = sq_lds_f2(tab_biased + off);
                const float fac = (mm * rinv2) * (rinv * fmaf(d, w.y, w.x));
                A.ax = fmaf(dx, fac, A.ax);
                A.ay = fmaf(dy, fac, A.ay);
                A.az = fmaf(dz, fac, A.az);
            }
        }
        // softened pairs (r < 2.8 eps), including the target meeting itself in its own leaf (r = 0)
        while(soft) {
            const int j = __ffs(soft) - 1;
            soft &= soft - 1;
            const float4 s = W.lst[4 * b + j];
            float dx = s.x - tx, dy = s.y - ty, dz = s.z - tz;
            if(WRAP) {
                dx -= P.f_boxs * rintf(dx * P.f_invboxs);
                dy -= P.f_boxs * rintf(dy * P.f_invboxs);
                dz -= P.f_boxs * rintf(dz * P.f_invboxs);
            }
            const float r = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
            const float2 sf = sq_soft_kernel(r, (double) P.f_hs);
            if(r < 511.f) {
                const float fl = floorf(r);
                const float dd = (r - fl) - 0.5f;
                const float4 w = sq_lds_f4(tab_biased + ((0x4B400000u + (unsigned) (int) fl) << 4));
                const float fac = s.w * sf.x * fmaf(dd, w.y, w.x);
                A.ax = fmaf(dx, fac, A.ax);
                A.ay = fmaf(dy, fac, A.ay);
                A.az = fmaf(dz, fac, A.az);
                if(POT) A.ap = fmaf(s.w * sf.y, fmaf(dd, w.w, w.z), A.ap);
            }
        }
    }
    A.dax += (double) A.ax; A.day += (double) A.ay; A.daz += (double) A.az;
    A.ax = A.ay = A.az = 0.f;
    if(POT) { A.dap += (double) A.ap; A.ap = 0.f; }
}

// Node constants of the FP32 tiers (one node per lane)
struct sq_node_f {
    float cmx, cmy, cmz, cx, cy, cz;
};

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

template <bool POT>
__global__ void __launch_bounds__(WALK_THREADS, 3)
sq_walk_kernel(const double4 * __restrict__ cm, const double4 * __restrict__ cl, const int4 * __restrict__ lk,
               const double * __restrict__ spos, const float * __restrict__ smass,
               const double * __restrict__ tpos, const float * __restrict__ tmass, const double * __restrict__ toldacc,
               const uint32_t * __restrict__ targets, long long ntargets, const float4 * __restrict__ gtab,
               double * __restrict__ accel, double * __restrict__ pot, unsigned long long * __restrict__ stats,
               int * __restrict__ group_counter, const sq_walk_par P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4 * tab = reinterpret_cast<float4 *>(smem_raw);
    sq_walk_warp_smem * Wall = reinterpret_cast<sq_walk_warp_smem *>(smem_raw + SQ_NTAB * sizeof(float4));
    for(int i = threadIdx.x; i < SQ_NTAB; i += blockDim.x)
        tab[i] = gtab[i];
    const unsigned lane = sq_lane();
    sq_walk_warp_smem & W = Wall[threadIdx.x >> 5];
    for(int i = lane; i < LIST_CAP; i += 32) {
        W.lst[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        W.lmask[i] = 0u;
    }
    __syncthreads();
    // table addresses biased so that the mantissa bits of (x + 1.5*2^23) index them directly
    const unsigned tab_biased = (unsigned) __cvta_generic_to_shared(tab) - (0x4B400000u << 4);

    const int G = P.gw;
    const int S = 32 / G;
    const int tl = lane & (G - 1), sl = lane / G;
    const long long ngroups = (ntargets + G - 1) / G;
    unsigned long long st_int = 0, st_ent = 0, st_tst = 0, st_pair = 0, st_exact = 0;

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
        const float tx = (float) ((px - gcx) * P.s), ty = (float) ((py - gcy) * P.s), tz = (float) ((pz - gcz) * P.s);
        __syncwarp();
        if(sl == 0) W.tgt[tl] = make_float4(tx, ty, tz, ta);
        // pairs may straddle the periodic image once group extent + table range reaches half the box
        const bool pairwrap = 2.f * ghm + P.f_range >= P.f_halfs;
        sq_walk_acc<POT> A;
        A.ax = A.ay = A.az = A.ap = 0.f;
        A.dax = A.day = A.daz = A.dap = 0.0;
        int top = 0, nlist = 0;
        if(lane == 0) W.stack[0] = make_int2(0, nt >= 32 ? (int) 0xffffffffu : (int) ((1u << nt) - 1u));
        top = 1;
        __syncwarp();

        while(top > 0) {
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
            float n_mass = 0.f;
            sq_node_f N;
            N.cmx = N.cmy = N.cmz = 0.f;
            if(have) {
                L = lk[node];
                const double4 M = cm[node];
                const double4 C = cl[node];
                st_tst++;
                const double dmx = sq_nearest(M.x - gcx, P.box, P.half), dmy = sq_nearest(M.y - gcy, P.box, P.half), dmz = sq_nearest(M.z - gcz, P.box, P.half);
                const double dcx = sq_nearest(C.x - gcx, P.box, P.half), dcy = sq_nearest(C.y - gcy, P.box, P.half), dcz = sq_nearest(C.z - gcz, P.box, P.half);
                N.cmx = (float) (dmx * P.s); N.cmy = (float) (dmy * P.s); N.cmz = (float) (dmz * P.s);
                N.cx = (float) (dcx * P.s); N.cy = (float) (dcy * P.s); N.cz = (float) (dcz * P.s);
                n_mass = (float) (M.w * P.minv);
                const float len = (float) (C.w * P.s);
                const float l2 = len * len, ml2 = n_mass * l2;
                const float acx = fabsf(N.cx), acy = fabsf(N.cy), acz = fabsf(N.cz);
                const float amx = fabsf(N.cmx), amy = fabsf(N.cmy), amz = fabsf(N.cmz);
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
                unsigned needk = 0, needo = 0, exact = 0;
                bool gopen = false;
                if(risky) exact = mask;
                else {
                    // ---- tier 1: the whole group at once (bounding box, min/max |a_old|)
                    const float qx = fmaxf(amx - ghx, 0.f), qy = fmaxf(amy - ghy, 0.f), qz = fmaxf(amz - ghz, 0.f);
                    const float wx = amx + ghx, wy = amy + ghy, wz = amz + ghz;
                    const float r2n = qx * qx + qy * qy + qz * qz, r2f = wx * wx + wy * wy + wz * wz;
                    const float en = fmaxf(fmaxf(acx - ghx, acy - ghy), acz - ghz);
                    const float ef = fmaxf(fmaxf(acx + ghx, acy + ghy), acz + ghz);
                    const float r2nl = fmaxf(r2n - dr2, 0.f), r2fh = r2f + dr2;
                    const bool gk_lo = (r2n <= rcA) || !(en > efA);
                    const bool gk_hi = (r2f <= rcB) || !(ef > efB);
                    const bool go_lo = (mlA > r2nl * r2nl * ta_lo) || (r2n < bhA) || (en < inA);
                    const bool go_hi = (mlB > r2fh * r2fh * ta_hi) || (r2f < bhB) || (ef < inB);
                    gopen = go_hi;
                    if(gk_lo) {
                        if(gk_hi) {
                            if(go_lo == go_hi) { if(go_hi) open_mask = mask; else acc_mask = mask; }
                            else needo = mask;
                        } else needk = mask;
                        if(go_lo != go_hi) needo = mask;
                    }
                }
                // ---- tier 2a: keep or discard, per target, FP32 with error intervals
                st_pair += __popc(needk) + __popc(needo);
                const bool omixed = needo != 0u;
                unsigned kept = (needk == 0u) ? needo : 0u;
                while(needk) {
                    const int t = 31 - __clz(needk);
                    const unsigned bit = 1u << t;
                    needk ^= bit;
                    const float4 T = W.tgt[t];
                    const float dx = N.cmx - T.x, dy = N.cmy - T.y, dz = N.cmz - T.z;
                    const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    const float em = fmaxf(fmaxf(fabsf(N.cx - T.x), fabsf(N.cy - T.y)), fabsf(N.cz - T.z));
                    const bool k_lo = (r2 <= rcA) || !(em > efA);
                    const bool k_hi = (r2 <= rcB) || !(em > efB);
                    if(k_lo != k_hi) exact |= bit;
                    else if(k_lo) kept |= bit;
                }
                if(!omixed) {
                    if(gopen) open_mask |= kept; else acc_mask |= kept;
                    kept = 0;
                }
                // ---- tier 2b: open or accept, for the kept targets of nodes whose opening is not uniform
                while(kept) {
                    const int t = 31 - __clz(kept);
                    const unsigned bit = 1u << t;
                    kept ^= bit;
                    const float4 T = W.tgt[t];
                    const float dx = N.cmx - T.x, dy = N.cmy - T.y, dz = N.cmz - T.z;
                    const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    const float em = fmaxf(fmaxf(fabsf(N.cx - T.x), fabsf(N.cy - T.y)), fabsf(N.cz - T.z));
                    const float r2l = fmaxf(r2 - dr2, 0.f), r2h = r2 + dr2;
                    const bool o_lo = (mlA > r2l * r2l * T.w) || (r2 < bhA) || (em < inA);
                    const bool o_hi = (mlB > r2h * r2h * T.w) || (r2 < bhB) || (em < inB);
                    if(o_lo != o_hi) exact |= bit;
                    else if(o_lo) open_mask |= bit;
                    else acc_mask |= bit;
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
            const int type = SQ_INFO_TYPE(L.w);
            // ---- accepted nodes -> interaction list
            {
                const unsigned b = __ballot_sync(SQ_FULL, acc_mask != 0);
                const int tot = __popc(b);
                if(tot) {
                    if(nlist + tot > LIST_CAP) {
                        __syncwarp();
                        st_ent += nlist;
                        if(pairwrap) sq_walk_flush<POT, true>(W, nlist, tab_biased, P, tx, ty, tz, tl, sl, S, A);
                        else sq_walk_flush<POT, false>(W, nlist, tab_biased, P, tx, ty, tz, tl, sl, S, A);
                        nlist = 0;
                        __syncwarp();
                    }
                    if(acc_mask) {
                        const int p = nlist + __popc(b & sq_lanemask_lt());
                        W.lst[p] = make_float4(N.cmx, N.cmy, N.cmz, n_mass);
                        W.lmask[p] = acc_mask;
                        st_int += __popc(acc_mask);
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
                    st_int += (unsigned long long) np * __popc(open_mask);
                    W.s_end[lane] = off + np;
                    W.s_first[lane] = L.x - off;        // particle = s_first + j
                    W.s_mask[lane] = open_mask;
                    __syncwarp();
                    for(int base = 0; base < tot; base += 32) {
                        const int chunk = (tot - base) < 32 ? (tot - base) : 32;
                        if(nlist + chunk > LIST_CAP) {
                            __syncwarp();
                            st_ent += nlist;
                            if(pairwrap) sq_walk_flush<POT, true>(W, nlist, tab_biased, P, tx, ty, tz, tl, sl, S, A);
                            else sq_walk_flush<POT, false>(W, nlist, tab_biased, P, tx, ty, tz, tl, sl, S, A);
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
                            W.lst[nlist + lane] = make_float4((float) (sq_nearest(sx - gcx, P.box, P.half) * P.s), (float) (sq_nearest(sy - gcy, P.box, P.half) * P.s),
                                                              (float) (sq_nearest(sz - gcz, P.box, P.half) * P.s), (float) ((double) smass[p] * P.minv));
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
        if(pairwrap) sq_walk_flush<POT, true>(W, nlist, tab_biased, P, tx, ty, tz, tl, sl, S, A);
        else sq_walk_flush<POT, false>(W, nlist, tab_biased, P, tx, ty, tz, tl, sl, S, A);
        // combine the slices of each target
        for(int o = G; o < 32; o <<= 1) {
            A.dax += __shfl_xor_sync(SQ_FULL, A.dax, o);
            A.day += __shfl_xor_sync(SQ_FULL, A.day, o);
            A.daz += __shfl_xor_sync(SQ_FULL, A.daz, o);
            if(POT) A.dap += __shfl_xor_sync(SQ_FULL, A.dap, o);
        }
        if(valid && sl == 0) {
            // gravshort2.hpp:88-107
            accel[3 * tidx] = A.dax * P.Ka;
            accel[3 * tidx + 1] = A.day * P.Ka;
            accel[3 * tidx + 2] = A.daz * P.Ka;
            if(POT) {
                const double m = (double) tmass[tidx];
                double p = A.dap * P.Kp;
                p += m / (P.h / 2.8);
                p -= 2.8372975 * pow(m, 2.0 / 3) * P.cbrtrho0;
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

// GravShortTable, libgadget/gravshort-tree2.cpp:29-69: float tables.  Device form: entry i holds
// (value at the middle of [i, i+1], slope), so w(r) = v + (r - i - 1/2) * slope; entry 511 = (0, 0):
// beyond the table the force is zero (gravity.h:53-54).
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
    float4 h[SQ_NTAB];     // (force value at mid-cell, force slope, potential value at mid-cell, potential slope)
    for(int i = 0; i < SQ_NTAB; i++) {
        if(i + 1 < SQ_NTAB) {
            const double sf = (double) ft[i + 1] - (double) ft[i], sp = (double) pt[i + 1] - (double) pt[i];
            h[i] = make_float4((float) ((double) ft[i] + 0.5 * sf), (float) sf, (float) ((double) pt[i] + 0.5 * sp), (float) sp);
        } else
            h[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    SQ_TRY(c->w_table.reserve(sizeof(h)));
    SQ_CUDA(cudaMemcpyAsync(c->w_table.p, h, sizeof(h), cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    return SQ_OK;
}

template <bool POT>
static int sq_walk_launch(sq_ctx * c, const uint32_t * targets, int64_t ntargets, const sq_walk_par & P)
{
    const size_t smem = SQ_NTAB * sizeof(float4) + WALK_WARPS * sizeof(sq_walk_warp_smem);
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
            c->w_table.as<float4>(), c->w_accel.as<double>(), c->w_pot.as<double>(), c->w_stats.as<unsigned long long>(),
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
