// Short-range tree walk: the interaction kernel.
//
// Replaces GravLocalTreeWalk::visit<PRIMARY> / apply_accn / shall_we_discard_node / shall_we_open_node
// (libgadget/gravshort2.hpp:152-358), TreeWalk::ev_primary (treewalk2.h:566-595), the GPU kernels
// treewalk_primary_kernel / treewalk_postprocess_kernel (treewalk2.cuh:104-173) and
// GravTreeOutput::postprocess (gravshort2.hpp:88-107).
//
// Design (not the reference's thread-per-particle depth-first walk):
//   * One warp walks the tree for a GROUP of G Peano-Hilbert-adjacent active particles (G = 8, 16 or 32) with ONE
//     node stack and ONE interaction list in shared memory.  A stack entry is (node, mask): mask = the targets of
//     the group that reach this node, i.e. opened every ancestor.  Per step the warp pops up to 32 entries, one
//     node per lane.  For the evaluation the warp is G targets x S = 32/G slices: slice s takes list entries
//     s, s+S, ...; smaller groups have a tighter bounding box, so fewer masked-off (target, source) slots.
//   * Coordinates are 32-bit fixed point (box / 2^32): node centres, centres of mass and particle positions are
//     quantised once per walk by sq_walk_prep_*; a group-relative coordinate is one integer subtraction (which
//     is also the periodic NEAREST()), one I2F and one multiply - no FP64 on the walk's hot path.  Node records
//     are two 128-bit loads + the link record; a source particle is one 128-bit load.
//   * The stack holds OPENED internal nodes (first child, child count, mask of the targets that opened it).  Per step the
//     warp pops as many entries as have <= 32 children in total and tests one child per lane, so the tests run at
//     close to full lane occupancy.
//   * Every (target, node) decision is the reference's own (discard: the Rcut + len/2 box; open: relative
//     criterion M l^2 > r^4 alpha |a_old|, Barnes-Hut angle, inside-0.6-len box), reached in tiers so that FP64 is
//     the exception:
//       1. group tier (FP32 interval arithmetic over the group's bounding box and its min/max |a_old|);
//       2. target tier (FP32, coordinates relative to the group centre in units of the split-table step, masses in
//          units of the mean particle mass): the class is evaluated at both ends of the error interval
//          (rounding + quantisation) of (r2, max-axis distance); equal -> decided;
//       3. exact tier: the reference's FP64 expressions on the original FP64 data with per-target NEAREST() for
//          the pairs whose interval straddles a threshold, and for nodes half a box away from the group.
//     The decision function is monotone in (r2, emax), so equal classes at the interval ends prove the FP64
//     class.  The interaction count therefore equals the reference's exactly.
//   * Evaluation is FP32: rsqrt, the erfc/"exact" split factor from a shared-memory table REPLICATED 16x so that
//     every lane of a half warp reads its own bank pair (conflict-free: 2.2 clk per lookup instead of 4.9,
//     measured by tools/ubench).  Distances are already in table units, so the table coordinate costs nothing.
//     Sources that may come within the softening length of a target (inside the group's box grown by h), and
//     everything when the periodic image can differ per pair (tiny boxes), go to a short "near" list evaluated
//     by a careful loop (spline softening gravshort2.hpp:332-350, self pair r = 0, per-pair wrap); the fast
//     loop has neither.  Partial sums are FP32 for one list flush (<= 128 entries), then FP64 accumulators.
//   * Epilogue = postprocess: x G, potential self terms; results written in sorted order.
#include "sq_common.cuh"
#include "shortrange_table.h"
#include <math.h>

#ifndef SQ_EVAL_UNROLL
#define SQ_EVAL_UNROLL 2
#endif
#ifndef SQ_PREFETCH
#define SQ_PREFETCH 0
#endif
#ifndef WALK_WARPS
#define WALK_WARPS 16
#endif
#define WALK_THREADS (WALK_WARPS * 32)
#ifndef STACK_CAP
#define STACK_CAP 384
#endif
#define STACK_RESERVE 160   /* 7 siblings x 22 levels: room a depth-first descent can always need */
#ifndef LIST_CAP
#define LIST_CAP 128        /* interaction-list entries per flush */
#endif
#define LIST_PAD 16         /* room to pad the last steps (4 steps x S entries) */
#ifndef NEAR_CAP
#define NEAR_CAP 32
#endif
#define TAB_ROWS (SQ_NTAB + 1)
#define TAB_COPIES 16
#define TAB_BYTES (TAB_ROWS * TAB_COPIES * 8)
#define SQ_EPS 4e-6f        /* relative slack on every FP32 threshold (>> 2^-24 rounding chains) */
#define SQ_FAR 3.0e4f       /* padding entries: far beyond the table, zero mask */
#define SQ_R2MAX 261888.f   /* 511.75^2: rows 511.. of the split table are zero */
#define SQ_MAGIC 12582912.f /* 1.5 * 2^23: adding it rounds to the nearest integer, kept in the mantissa */

struct sq_walk_par {
    // exact tier (reference expressions, FP64)
    double box, half, rcut, rcut2, bh2, h, errtol;
    // scaling: lengths x s (table units), masses x minv
    double s, minv, Ka, Kp;       // accel = Ka * sum, potential = Kp * sum (+ self terms)
    double G, cbrtrho0;
    double qinv;                  // 2^32 / box: FP64 position -> fixed point
    int relative;                 // TreeUseBH == 0
    // FP32 tiers (table units)
    float qs;                     // table units per fixed-point unit
    float f_rcut, f_rc2_lo, f_rc2_hi, f_invbh2, f_h2, f_hs, f_boxs, f_invboxs, f_halfs, f_range;
    // exact keep test per tree level: |centre - x| <= e_in certainly inside the keep box, > e_out certainly outside
    // (fixed-point quanta); e_in < 0: -1 = every target is inside (Rcut + len/2 >= Box/2), -2 = use the FP32/FP64 tiers
    int e_in[SQ_MAXLEVEL + 3], e_out[SQ_MAXLEVEL + 3];
    // children of a level-L node sit at centre +- q4[L] (len / 4 in quanta, rounded down); e_child[L] = half size of a
    // child's keep box + slack (e_out[L + 1] + 8), or -1 when no child can be ruled out at push time
    int q4[SQ_MAXLEVEL + 3], e_child[SQ_MAXLEVEL + 3];
};

struct sq_walk_warp_smem {
    // interaction list, structure of arrays so that one 128-bit LDS brings the same coordinate of 4 consecutive steps
    // of a slice into an aligned register quad (= two operands of the packed FP32 instructions): the entry at list
    // position p (step i = p / S, slice p % S) lives at sq_list_slot<S>(p)
    float lx[LIST_CAP + LIST_PAD], ly[LIST_CAP + LIST_PAD], lz[LIST_CAP + LIST_PAD], lm[LIST_CAP + LIST_PAD];
    float4 nl[NEAR_CAP];               // near list (careful loop)
    int2 stack[STACK_CAP];
    unsigned mw[LIST_CAP + LIST_PAD];  // mask words, one per evaluation step: bit (s*G + t) = entry of slice s reaches target t
    unsigned nm[NEAR_CAP];
    float4 nrec[32][3];                // tier 2 work list: (cm.xyz, dr2) (centre.xyz, de) (len, M l^2, need mask, -)
    int s_first[32];                     // pop: children of the popped entries; later in the step: particles of opened leaves
    unsigned s_mask[32];
    int sx[3][32];                     // target coordinates per axis, sorted (fixed point, relative to the first target)
    unsigned pmk[3][36];               // pmk[a][k] = targets among the k smallest coordinates of axis a
    unsigned char scnt[STACK_CAP];     // stack entry i: octants of its children that survived the push-time box test
    unsigned char scm[STACK_CAP];      // stack entry i: octants that have a child (child k = k-th set bit)
    unsigned char src[256];            // slot j of an expansion (children of popped entries / particles of opened leaves) -> lane that owns it
};

static_assert(TAB_BYTES + WALK_WARPS * sizeof(sq_walk_warp_smem) <= 232448, "walk kernel shared memory exceeds the 227 KB of an sm_100 CTA");

// FP64 position -> 32-bit fixed point, periodic (two's complement wrap = the box)
__host__ __device__ __forceinline__ unsigned sq_quant(double x, double qinv)
{
    return (unsigned) (long long) floor(x * qinv);
}

__device__ __forceinline__ float2 sq_lds_f2(unsigned addr)
{
    float2 v;       // only used on the split table, constant after the kernel prologue
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

__device__ __forceinline__ float sq_sqrt(float x)
{
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
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


__device__ __forceinline__ int sq_add_sat(int a, int b)
{
    int r;
    asm("add.sat.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ int sq_sub_sat(int a, int b)
{
    int r;
    asm("sub.sat.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}
// number of the 32 sorted values that are <= v
__device__ __forceinline__ int sq_count_le(const int * s, int v)
{
    int pos = 0;
    #pragma unroll
    for(int step = 16; step >= 1; step >>= 1)
        if(s[pos + step - 1] <= v) pos += step;
    if(s[pos] <= v) pos++;
    return pos;
}
// targets with |c - x_t| <= e_in on this axis; amb is set when some target sits within (e_in, e_out] of c
__device__ __forceinline__ unsigned sq_axis_keep(const int * s, const unsigned * pm, int c, int e_in, int e_out, bool & amb)
{
    const int hi = sq_add_sat(c, e_in), lo = sq_sub_sat(c, e_in);
    const int p = sq_count_le(s, hi);                              // values <= c + e_in
    const int q = lo == (int) 0x80000000 ? 0 : sq_count_le(s, lo - 1);   // values < c - e_in
    if(p < 32 && s[p] != 0x7fffffff && s[p] <= sq_add_sat(c, e_out)) amb = true;
    if(q > 0 && s[q - 1] >= sq_sub_sat(c, e_out)) amb = true;
    return pm[p] & ~pm[q];
}

template <int S>
__device__ __forceinline__ int sq_list_slot(int p)
{
    const int i = p / S, sl = p % S;
    return (((i >> 2) * S + sl) << 2) + (i & 3);
}

// ---- fast evaluation: lane = (slice s, target t); step i gives slice s the entry i*S + s and mask word i, whose bit
// `lane` says whether that entry reaches this lane's target.  Two steps at a time in packed FP32 (FADD2 / FMUL2 /
// FFMA2 of sm_100: two results per issue slot; MUFU, the table reads and the mask selects stay scalar).
// Every genuine entry is farther than the softening length from every target of the group (the others are in the
// near list), padding entries are SQ_FAR away, so r2 > 0 and no softening or wrap test is needed.
// Table (gravity.h:50-60): i = floor(r), linear interpolation, zero from row 511 on.
//   !POT: row i = (intercept, slope) of the line through table points i and i+1 (value = intercept + r * slope: no
//         fractional part needed), 16 copies of 8 bytes per 128-byte row
//    POT: row i = (force intercept, force slope, potential intercept, potential slope), 8 copies of 16 bytes per row: a
//         128-bit LDS is served 8 lanes at a time, each lane of a quarter warp has its own copy (conflict-free)
//   rows 511.. = 0: the clamp of r2 to 511.75^2 lands there, which is the reference's "outside the table"
struct sq_acc2 { float2 x, y, z, p; };

template <bool POT>
__device__ __forceinline__ void sq_eval_pair(float x0, float x1, float y0, float y1, float z0, float z1, float m0, float m1, bool on0, bool on1,
                                             unsigned tabaddr, float2 ntx, float2 nty, float2 ntz, sq_acc2 & A)
{
    const float2 dx = __fadd2_rn(make_float2(x0, x1), ntx), dy = __fadd2_rn(make_float2(y0, y1), nty), dz = __fadd2_rn(make_float2(z0, z1), ntz);
    float2 r2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
    // beyond the table every row is zero: clamping r2 (and with it rinv) only keeps the row index in range
    r2.x = fminf(r2.x, SQ_R2MAX); r2.y = fminf(r2.y, SQ_R2MAX);
    const float2 rinv = make_float2(sq_rsqrt(r2.x), sq_rsqrt(r2.y));
    // r for the interpolation; floor(r) straight into the mantissa by a round-down FMA onto 1.5 * 2^23
    const float2 r = __fmul2_rn(r2, rinv);
    const float2 tt = __ffma2_rd(r2, rinv, make_float2(SQ_MAGIC, SQ_MAGIC));
    const unsigned a0 = ((unsigned) __float_as_int(tt.x) << 7) + tabaddr, a1 = ((unsigned) __float_as_int(tt.y) << 7) + tabaddr;
    const float2 rinv2 = __fmul2_rn(rinv, rinv);
    if(POT) {
        const float4 e0 = sq_lds_f4(a0), e1 = sq_lds_f4(a1);
        // scalar FMAs here: pairing (e0.y, e1.y) etc. for the packed form costs more register moves than it saves
        const float2 wf = make_float2(fmaf(r.x, e0.y, e0.x), fmaf(r.y, e1.y, e1.x));
        const float2 wp = make_float2(fmaf(r.x, e0.w, e0.z), fmaf(r.y, e1.w, e1.z));
        const float2 mm = make_float2(on0 ? m0 : 0.f, on1 ? m1 : 0.f);
        const float2 mr = __fmul2_rn(mm, rinv);
        const float2 fac = __fmul2_rn(__fmul2_rn(mr, rinv2), wf);
        A.x = __ffma2_rn(dx, fac, A.x);
        A.y = __ffma2_rn(dy, fac, A.y);
        A.z = __ffma2_rn(dz, fac, A.z);
        A.p = __ffma2_rn(mr, wp, A.p);       // sign applied at the flush
    } else {
        const float2 w0 = sq_lds_f2(a0), w1 = sq_lds_f2(a1);
        const float2 wf = make_float2(fmaf(r.x, w0.y, w0.x), fmaf(r.y, w1.y, w1.x));
        const float2 mm = make_float2(on0 ? m0 : 0.f, on1 ? m1 : 0.f);
        const float2 fac = __fmul2_rn(__fmul2_rn(mm, rinv2), __fmul2_rn(rinv, wf));
        A.x = __ffma2_rn(dx, fac, A.x);
        A.y = __ffma2_rn(dy, fac, A.y);
        A.z = __ffma2_rn(dz, fac, A.z);
    }
}

template <bool POT, int S>
__device__ __forceinline__ void sq_eval_list(const sq_walk_warp_smem & W, int nsteps4, unsigned tabaddr, float tx, float ty, float tz,
                                             int s, unsigned lanebit, float & ax, float & ay, float & az, float & ap)
{
    const uint4 * mw4 = reinterpret_cast<const uint4 *>(W.mw);
    const float4 * X = reinterpret_cast<const float4 *>(W.lx) + s, * Y = reinterpret_cast<const float4 *>(W.ly) + s;
    const float4 * Z = reinterpret_cast<const float4 *>(W.lz) + s, * M = reinterpret_cast<const float4 *>(W.lm) + s;
    const float2 ntx = make_float2(-tx, -tx), nty = make_float2(-ty, -ty), ntz = make_float2(-tz, -tz);
    sq_acc2 A;
    A.x = A.y = A.z = A.p = make_float2(0.f, 0.f);
    const int nb = nsteps4 >> 2;
    constexpr int eval_unroll = SQ_EVAL_UNROLL;
    #pragma unroll eval_unroll
    for(int b = 0; b < nb; b++) {
        const uint4 w4 = mw4[b];
        const float4 x4 = X[b * S], y4 = Y[b * S], z4 = Z[b * S], m4 = M[b * S];
        sq_eval_pair<POT>(x4.x, x4.y, y4.x, y4.y, z4.x, z4.y, m4.x, m4.y, (w4.x & lanebit) != 0u, (w4.y & lanebit) != 0u, tabaddr, ntx, nty, ntz, A);
        sq_eval_pair<POT>(x4.z, x4.w, y4.z, y4.w, z4.z, z4.w, m4.z, m4.w, (w4.z & lanebit) != 0u, (w4.w & lanebit) != 0u, tabaddr, ntx, nty, ntz, A);
    }
    ax += A.x.x + A.x.y; ay += A.y.x + A.y.y; az += A.z.x + A.z.y;
    if(POT) ap -= A.p.x + A.p.y;
}

// ---- careful evaluation of the near list (slice s takes entries s, s+S, ...): softened pairs (r < 2.8 eps),
// the target meeting itself in its own leaf (r = 0, as in the reference), per-pair periodic wrap.
template <bool POT>
__device__ __forceinline__ void sq_near_pair(const float4 src, unsigned tabrow0, const sq_walk_par & P, bool pairwrap, bool soft_only,
                                             float tx, float ty, float tz, float & ax, float & ay, float & az, float & ap, bool & deferred)
{
    float dx = src.x - tx, dy = src.y - ty, dz = src.z - tz;
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
        if(!soft_only) { deferred = true; return; }
        r = sqrtf(r2);
        const float2 sf = sq_soft_kernel(r, P.f_hs);
        fac = sf.x; facpot = sf.y;
    }
    // i = floor(r) clamped to the zero rows from 511 on
    const float fl = fminf(floorf(r), 511.f);
    const unsigned addr = tabrow0 + ((unsigned) (int) fl << 7);
    float wf, wp = 0.f;
    if(POT) {
        const float4 e = sq_lds_f4(addr);
        wf = fmaf(r, e.y, e.x); wp = fmaf(r, e.w, e.z);
    } else {
        const float2 w = sq_lds_f2(addr);
        wf = fmaf(r, w.y, w.x);
    }
    const float f = src.w * fac * wf;
    ax = fmaf(dx, f, ax);
    ay = fmaf(dy, f, ay);
    az = fmaf(dz, f, az);
    if(POT) ap = fmaf(src.w * facpot, wp, ap);
}

// ---- careful evaluation of the near list (slice s takes entries s, s+S, ...): softened pairs (r < 2.8 eps),
// the target meeting itself in its own leaf (r = 0, as in the reference), per-pair periodic wrap.
// Softened pairs are rare per lane (typically the self pair only) but nearly every entry has one in SOME lane, so a
// lane queues its softened entries (two slots) and the FP64 spline runs once or twice per flush instead of per entry.
template <bool POT, int S>
__device__ __forceinline__ void sq_eval_near(const sq_walk_warp_smem & W, int nn, unsigned tabrow0, const sq_walk_par & P, bool pairwrap,
                                             float tx, float ty, float tz, int s, int t, float & ax, float & ay, float & az, float & ap)
{
    int nq = 0, q0 = 0, q1 = 0;
    for(int k = s; k < nn; k += S) {
        const float4 src = W.nl[k];
        const unsigned m = W.nm[k];
        if((m >> t) & 1u) {
            bool deferred = false;
            sq_near_pair<POT>(src, tabrow0, P, pairwrap, false, tx, ty, tz, ax, ay, az, ap, deferred);
            if(deferred) {
                if(nq == 0) q0 = k;
                else if(nq == 1) q1 = k;
                else sq_near_pair<POT>(src, tabrow0, P, pairwrap, true, tx, ty, tz, ax, ay, az, ap, deferred);
                nq++;
            }
        }
    }
    #pragma unroll 1
    for(int j = 0; j < 2; j++) {
        if(!__any_sync(SQ_FULL, nq > j)) break;
        if(nq > j) {
            bool deferred = false;
            sq_near_pair<POT>(W.nl[j == 0 ? q0 : q1], tabrow0, P, pairwrap, true, tx, ty, tz, ax, ay, az, ap, deferred);
        }
    }
}

// mask field of list entry p inside the per-step mask words (G bits per entry, S entries per word)
template <int G>
__device__ __forceinline__ void sq_store_field(unsigned * mw, int p, unsigned m)
{
    if(G == 8) reinterpret_cast<unsigned char *>(mw)[p] = (unsigned char) m;
    else if(G == 16) reinterpret_cast<unsigned short *>(mw)[p] = (unsigned short) m;
    else mw[p] = m;
}

template <bool POT, int G>
__global__ void __launch_bounds__(WALK_THREADS, 1)
sq_walk_kernel(const uint4 * __restrict__ icm, const uint4 * __restrict__ icl, const int4 * __restrict__ lk,
               const double4 * __restrict__ cm, const double4 * __restrict__ cl, const uint4 * __restrict__ ipart,
               const double * __restrict__ tpos, const float * __restrict__ tmass, const double * __restrict__ toldacc,
               const uint32_t * __restrict__ targets, long long ntargets, const float2 * __restrict__ gtab,
               double * __restrict__ accel, double * __restrict__ pot, unsigned long long * __restrict__ stats,
               int * __restrict__ group_counter, unsigned * __restrict__ gcost, const sq_walk_par P)
{
    constexpr int S = 32 / G;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2 * tab = reinterpret_cast<float2 *>(smem_raw);
    sq_walk_warp_smem * Wall = reinterpret_cast<sq_walk_warp_smem *>(smem_raw + TAB_BYTES);
    // copies of every table row side by side in a 128-byte row: 16 x float2 (force only) or 8 x float4 (force + potential)
    if(POT) {
        const float4 * g4 = reinterpret_cast<const float4 *>(gtab);
        float4 * tab4 = reinterpret_cast<float4 *>(smem_raw);
        for(int i = threadIdx.x; i < TAB_ROWS * (TAB_COPIES / 2); i += blockDim.x)
            tab4[i] = g4[i / (TAB_COPIES / 2)];
    } else {
        for(int i = threadIdx.x; i < TAB_ROWS * TAB_COPIES; i += blockDim.x)
            tab[i] = gtab[i / TAB_COPIES];
    }
    const unsigned lane = sq_lane();
    sq_walk_warp_smem & W = Wall[threadIdx.x >> 5];
    __syncthreads();
    const unsigned tabrow0 = (unsigned) __cvta_generic_to_shared(tab) + (POT ? (lane & 7u) * 16u : (lane & 15u) * 8u);
    // biased so that the mantissa bits of (x + 1.5*2^23) index the rows directly
    const unsigned tabaddr = tabrow0 - (0x4B400000u << 7);
    const int t = (int) (lane & (G - 1)), s = (int) (lane / G);
    const unsigned lanebit = 1u << lane;
    const float INF = __int_as_float(0x7f800000);

    const long long ngroups = (ntargets + G - 1) / G;
    unsigned long long st_int = 0, st_ent = 0, st_tst = 0, st_pair = 0, st_exact = 0, st_pseudo = 0;

    for(;;) {
        long long g = 0;
        if(lane == 0) g = atomicAdd(group_counter, 1);
        g = __shfl_sync(SQ_FULL, g, 0);
        if(g >= ngroups) break;
        const long long t0 = g * G;
        const unsigned long long st_int_before = st_int;
        const int nt = (int) ((ntargets - t0) < G ? (ntargets - t0) : G);
        const bool valid = t < nt;
        long long tidx = 0;
        unsigned ix = 0, iy = 0, iz = 0;
        float ta = INF;
        if(valid) {
            tidx = targets ? (long long) targets[t0 + t] : (t0 + t);
            ix = sq_quant(tpos[3 * tidx], P.qinv); iy = sq_quant(tpos[3 * tidx + 1], P.qinv); iz = sq_quant(tpos[3 * tidx + 2], P.qinv);
            // errtol*|a_old| in the scaled system: M l^2 > r^4 a  <=>  M' l'^2 > r'^4 a / (munit s^2); BH mode: never
            if(P.relative) ta = (float) (P.errtol * toldacc[tidx] * (P.minv / (P.s * P.s)));
        }
        // 1 / sqrt(errtol |a_old|) for the per-target relative criterion (0: never opens - Barnes-Hut mode, idle lanes)
        const float rta = (valid && P.relative) ? sq_rsqrt(ta) : 0.f;
        // group bounding box in fixed point, relative to the first target (always valid)
        const unsigned ix0 = __shfl_sync(SQ_FULL, ix, 0), iy0 = __shfl_sync(SQ_FULL, iy, 0), iz0 = __shfl_sync(SQ_FULL, iz, 0);
        const int rx = valid ? (int) (ix - ix0) : 0, ry = valid ? (int) (iy - iy0) : 0, rz = valid ? (int) (iz - iz0) : 0;
        int mnx = rx, mxx = rx, mny = ry, mxy = ry, mnz = rz, mxz = rz;
        float ta_lo = valid ? ta : INF, ta_hi = valid ? ta : 0.f;
        #pragma unroll
        for(int o = G / 2; o >= 1; o >>= 1) {
            mnx = min(mnx, __shfl_xor_sync(SQ_FULL, mnx, o)); mxx = max(mxx, __shfl_xor_sync(SQ_FULL, mxx, o));
            mny = min(mny, __shfl_xor_sync(SQ_FULL, mny, o)); mxy = max(mxy, __shfl_xor_sync(SQ_FULL, mxy, o));
            mnz = min(mnz, __shfl_xor_sync(SQ_FULL, mnz, o)); mxz = max(mxz, __shfl_xor_sync(SQ_FULL, mxz, o));
            ta_lo = fminf(ta_lo, __shfl_xor_sync(SQ_FULL, ta_lo, o)); ta_hi = fmaxf(ta_hi, __shfl_xor_sync(SQ_FULL, ta_hi, o));
        }
        ta_lo *= (1.f - SQ_EPS);
        ta_hi = P.relative ? ta_hi * (1.f + SQ_EPS) : INF;
        // extents as unsigned differences: a group may span more than half the box (then everything is "risky")
        const unsigned exx = (unsigned) mxx - (unsigned) mnx, exy = (unsigned) mxy - (unsigned) mny, exz = (unsigned) mxz - (unsigned) mnz;
        const unsigned gcix = ix0 + (unsigned) mnx + (exx >> 1), gciy = iy0 + (unsigned) mny + (exy >> 1), gciz = iz0 + (unsigned) mnz + (exz >> 1);
        // half extents in table units, rounded up
        const float ghx = (float) exx * (0.5f * 1.000001f) * P.qs + 2.f * P.qs;
        const float ghy = (float) exy * (0.5f * 1.000001f) * P.qs + 2.f * P.qs;
        const float ghz = (float) exz * (0.5f * 1.000001f) * P.qs + 2.f * P.qs;
        const float ghm = fmaxf(ghx, fmaxf(ghy, ghz));
        // idle lanes sit far away so that they never see r = 0
        const float tx = valid ? (float) (int) (ix - gcix) * P.qs : 2.f * SQ_FAR;
        const float ty = valid ? (float) (int) (iy - gciy) * P.qs : 2.f * SQ_FAR;
        const float tz = valid ? (float) (int) (iz - gciz) * P.qs : 2.f * SQ_FAR;
        __syncwarp();
        // sources inside this box may come within the softening length of a target -> near list
        const float nhx = ghx + P.f_hs * 1.001f + 1e-3f, nhy = ghy + P.f_hs * 1.001f + 1e-3f, nhz = ghz + P.f_hs * 1.001f + 1e-3f;
        // pairs may straddle the periodic image once group extent + table range reaches half the box
        const bool pairwrap = 2.f * ghm + P.f_range >= P.f_halfs;
        float ax = 0.f, ay = 0.f, az = 0.f, ap = 0.f;
        double dax = 0.0, day = 0.0, daz = 0.0, dap = 0.0;
        int top = 1, nlist = 0, nn = 0;
        // ---- exact keep test: sort the targets' coordinates on every axis, prefix masks over the sorted order
        // (a group wider than a quarter of the box keeps the FP32 / FP64 tiers only)
        const bool intkeep = max(exx, max(exy, exz)) < 0x40000000u;
        if(intkeep) {
            unsigned * oh = reinterpret_cast<unsigned *>(W.nrec);
            #pragma unroll
            for(int a = 0; a < 3; a++) {
                const int v = valid ? (a == 0 ? rx : (a == 1 ? ry : rz)) : 0x7fffffff;
                int rk = 0;
                #pragma unroll 8
                for(int u = 0; u < 32; u++) {
                    const int vu = __shfl_sync(SQ_FULL, v, u);
                    rk += (vu < v || (vu == v && u < (int) lane)) ? 1 : 0;
                }
                W.sx[a][rk] = v;
                oh[a * 32 + rk] = valid ? lanebit : 0u;
            }
            __syncwarp();
            #pragma unroll
            for(int a = 0; a < 3; a++) {
                unsigned m = oh[a * 32 + lane];
                #pragma unroll
                for(int o = 1; o < 32; o <<= 1) {
                    const unsigned y = __shfl_up_sync(SQ_FULL, m, o);
                    if((int) lane >= o) m |= y;
                }
                W.pmk[a][lane + 1] = m;
                if(lane == 0) W.pmk[a][0] = 0u;
            }
        }
        if(lane == 0) {
            // the root is "the only child" of a virtual entry reached by every target
            W.stack[0] = make_int2(0, nt >= 32 ? (int) 0xffffffffu : (int) ((1u << nt) - 1u));
            W.scnt[0] = 1; W.scm[0] = 1;
        }
        __syncwarp();

        for(;;) {
            // the last pass (empty stack) only drains the lists
            const bool drain = top == 0;
            unsigned acc_mask = 0, open_mask = 0;
            int4 L = make_int4(0, 0, -1, 0);
            float4 nodeE = make_float4(0.f, 0.f, 0.f, 0.f);
            uint4 nodeC = make_uint4(0u, 0u, 0u, 0u);
            int node = 0;
            unsigned t2_need = 0, risky_exact = 0;
            bool relonly = false;
            float4 t2_cm = nodeE, t2_cl = nodeE, t2_lm = nodeE;
            if(!drain) {
                // pop the entries whose children fit the 32 lanes (always at least one: <= 8 children); every child can
                // push one entry, so wide pops only while a depth-first descent (<= 7 entries per level) still fits
                const bool ontop = (int) lane < top;
                const int2 ent = ontop ? W.stack[top - 1 - lane] : make_int2(0, 0);
                const unsigned sv = ontop ? (unsigned) W.scnt[top - 1 - lane] : 0u;
                const unsigned cmk = ontop ? (unsigned) W.scm[top - 1 - lane] : 0u;
                const int cnt = ontop ? __popc(sv) : 64;
                int tot_all;
                const int exc = sq_warp_excl_scan(cnt, tot_all);
                int npop = __popc(__ballot_sync(SQ_FULL, exc + cnt <= 32));
                if(top - npop + 32 > STACK_CAP - STACK_RESERVE) npop = 1;
                const int nchild = __shfl_sync(SQ_FULL, exc + cnt, npop - 1);
                // slot j of the expansion belongs to the popped entry whose first slot is the last start marker <= j
                const unsigned starts = __reduce_or_sync(SQ_FULL, (int) lane < npop ? (1u << (exc & 31)) : 0u);
                const bool have = (int) lane < nchild;
                const int el = have ? __popc(starts & ((2u << lane) - 1u)) - 1 : 0;
                const int e_first = __shfl_sync(SQ_FULL, ent.x, el);
                const unsigned e_mask = (unsigned) __shfl_sync(SQ_FULL, ent.y, el);
                const unsigned e_pack = __shfl_sync(SQ_FULL, sv | (cmk << 8) | ((unsigned) exc << 16), el);
                top -= npop;
                unsigned mask = 0;
                if(have) {
                    // the (lane - first slot)-th surviving octant of that entry -> its position among the children
                    unsigned m8 = e_pack & 0xffu;
                    int nth = (int) lane - (int) (e_pack >> 16), pos = 0;
                    int c = __popc(m8 & 0xfu);
                    if(nth >= c) { nth -= c; pos = 4; m8 >>= 4; }
                    c = __popc(m8 & 3u);
                    if(nth >= c) { nth -= c; pos += 2; m8 >>= 2; }
                    if(nth >= (int) (m8 & 1u)) pos += 1;
                    node = e_first + __popc((e_pack >> 8) & 0xffu & ((1u << pos) - 1u));
                    mask = e_mask;
                }
                __syncwarp();
                if(have) {
                    L = lk[node];
                    const uint4 A = icm[node], B = icl[node];
                    nodeC = B;
                    st_tst++;
                    const float cmx = (float) (int) (A.x - gcix) * P.qs, cmy = (float) (int) (A.y - gciy) * P.qs, cmz = (float) (int) (A.z - gciz) * P.qs;
                    const float cx = (float) (int) (B.x - gcix) * P.qs, cy = (float) (int) (B.y - gciy) * P.qs, cz = (float) (int) (B.z - gciz) * P.qs;
                    const float n_mass = __uint_as_float(A.w), len = __uint_as_float(B.w);
                    nodeE = make_float4(cmx, cmy, cmz, n_mass);
                    const float l2 = len * len, ml2 = n_mass * l2;
                    const float acx = fabsf(cx), acy = fabsf(cy), acz = fabsf(cz);
                    const float amx = fabsf(cmx), amy = fabsf(cmy), amz = fabsf(cmz);
                    // error bounds of the FP32 coordinates (rounding + fixed-point quantisation) / squared distances
                    const float scale = fmaxf(fmaxf(fmaxf(acx, acy), fmaxf(acz, amx)), fmaxf(amy, amz)) + ghm;
                    const float de = fmaf(3e-7f, scale, 2.5f * P.qs);
                    const float R1 = amx + amy + amz + 3.f * ghm;
                    const float dr2 = 4.f * R1 * de + 1e-6f * R1 * R1;
                    const float eff = fmaf(0.5f, len, P.f_rcut), ins = 0.6f * len;
                    // thresholds with the relative slack and the coordinate error folded in:
                    //   "A" = the end of the interval that keeps / opens most, "B" = the end that keeps / opens least
                    const float rcA = P.f_rc2_hi + dr2, rcB = P.f_rc2_lo - dr2;
                    const float efA = fmaf(eff, 1.f + SQ_EPS, de), efB = fmaf(eff, 1.f - SQ_EPS, -de);
                    const float inA = fmaf(ins, 1.f + SQ_EPS, de), inB = fmaf(ins, 1.f - SQ_EPS, -de);
                    const float bhq = l2 * P.f_invbh2;
                    const float bhA = fmaf(bhq, 1.f + SQ_EPS, dr2), bhB = fmaf(bhq, 1.f - SQ_EPS, -dr2);
                    const float mlA = ml2 * (1.f + SQ_EPS), mlB = ml2 * (1.f - SQ_EPS);
                    // a node about half a box away from the group: per-target NEAREST may differ from the group's
                    const bool risky = fmaxf(fmaxf(acx + ghx, acy + ghy), fmaxf(acz + ghz, fmaxf(fmaxf(amx + ghx, amy + ghy), amz + ghz))) >= P.f_halfs * 0.999f;
                    unsigned need = 0;
                    if(risky) risky_exact = mask;
                    else {
                        // With the centre of mass inside the cell "r2 <= Rcut^2" implies the box test, so the box
                        // decides keep/discard.
                        const float en = fmaxf(fmaxf(acx - ghx, acy - ghy), acz - ghz);
                        const float ef = fmaxf(fmaxf(acx + ghx, acy + ghy), acz + ghz);
                        const float hl = fmaf(0.5f * (1.f + 1e-5f), len, de);
                        const bool cmin = fabsf(cmx - cx) <= hl && fabsf(cmy - cy) <= hl && fabsf(cmz - cz) <= hl;
                        const int lvl = SQ_INFO_LEVEL(L.w);
                        const int e_in = P.e_in[lvl], e_out = P.e_out[lvl];
                        if(!cmin) need = mask;
                        else if(en > efA) { }                       // outside every target's box: discarded
                        else {
                            // ---- tier 0: the targets that keep the node, exact
                            unsigned keep = mask;
                            bool known = !(ef > efB);               // inside every target's box
                            const bool known_exact = known || (intkeep && e_in != -2);
                            if(!known && intkeep && e_in != -2) {
                                known = true;
                                if(e_in >= 0) {
                                    // only the faces of the keep box that cut through the group's extent need a search
                                    const int c0 = (int) (B.x - ix0), c1 = (int) (B.y - iy0), c2 = (int) (B.z - iz0);
                                    unsigned tasks = 0;
                                    if(sq_add_sat(c0, e_in) < mxx) tasks |= 1u;
                                    if(sq_sub_sat(c0, e_in) > mnx) tasks |= 2u;
                                    if(sq_add_sat(c1, e_in) < mxy) tasks |= 4u;
                                    if(sq_sub_sat(c1, e_in) > mny) tasks |= 8u;
                                    if(sq_add_sat(c2, e_in) < mxz) tasks |= 16u;
                                    if(sq_sub_sat(c2, e_in) > mnz) tasks |= 32u;
                                    bool amb = false;
                                    while(tasks) {
                                        const int j = __ffs((int) tasks) - 1;
                                        tasks &= tasks - 1u;
                                        const int a = j >> 1;
                                        const bool lower = (j & 1) != 0;
                                        const int c = a == 0 ? c0 : (a == 1 ? c1 : c2);
                                        const int * sa = W.sx[a];
                                        // upper face: targets with x <= c + e_in pass; lower face: targets with x < c - e_in fail
                                        const int v = lower ? sq_sub_sat(c, e_in) - 1 : sq_add_sat(c, e_in);     // the task bit implies no saturation at INT_MIN
                                        const int k = sq_count_le(sa, v);
                                        const unsigned pm = W.pmk[a][k];
                                        if(lower) {
                                            if(k > 0 && sa[k - 1] >= sq_sub_sat(c, e_out)) amb = true;
                                            keep &= ~pm;
                                        } else {
                                            if(k < 32 && sa[k] != 0x7fffffff && sa[k] <= sq_add_sat(c, e_out)) amb = true;
                                            keep &= pm;
                                        }
                                    }
                                    if(amb) { risky_exact = mask; keep = 0u; }
                                }
                            }
                            if(!known) need = mask;
                            else if(keep == 0u) { }
                            else if(SQ_INFO_TYPE(L.w) == 0 && L.y == 1) acc_mask = keep;   // one particle: monopole = the particle
                            else {
                                // ---- tier 1: one opening decision for all kept targets (group box clipped to the keep box)
                                const float qx = fmaxf(amx - ghx, 0.f), qy = fmaxf(amy - ghy, 0.f), qz = fmaxf(amz - ghz, 0.f);
                                const float wx = fminf(amx + ghx, fabsf(cmx - cx) + efA), wy = fminf(amy + ghy, fabsf(cmy - cy) + efA), wz = fminf(amz + ghz, fabsf(cmz - cz) + efA);
                                const float r2n = fmaf(qz, qz, fmaf(qy, qy, qx * qx)), r2f = fmaf(wz, wz, fmaf(wy, wy, wx * wx));
                                const float r2nl = fmaxf(r2n - dr2, 0.f), r2fh = r2f + dr2;
                                const bool go_lo = (mlA > r2nl * r2nl * ta_lo) || (r2n < bhA) || (en < inA);
                                const bool go_hi = (mlB > r2fh * r2fh * ta_hi) || (r2f < bhB) || (ef < inB);
                                if(go_lo == go_hi) {
                                    if(go_hi) open_mask = keep; else acc_mask = keep;
                                } else {
                                    need = keep;
                                    // kept for certain, and neither the angle nor the inside test can fire for any target:
                                    // only the relative criterion is left to tier 2
                                    relonly = known_exact && !(r2n < bhA) && !(en < inA);
                                }
                            }
                        }
                        if(need) {
                            // ---- tier 2 runs transposed below: hand the node over
                            t2_need = need;
                            t2_cm = make_float4(cmx, cmy, cmz, dr2);
                            t2_cl = make_float4(cx, cy, cz, de);
                            t2_lm = make_float4(len, sq_sqrt(ml2), 0.f, relonly ? 1.f : 0.f);
                        }
                    }
                }
            }
            // ---- tier 2, transposed: lane = (slice s, target t); slice s classifies one handed-over node per step for
            // all G targets at once (FP32 with error intervals), the masks come back through ballots
            {
                const unsigned fb = __ballot_sync(SQ_FULL, t2_need != 0u);
                if(fb) {
                    const int nf = __popc(fb);
                    const int slot = __popc(fb & sq_lanemask_lt());
                    if(t2_need) {
                        t2_lm.z = __uint_as_float(t2_need);
                        W.nrec[slot][0] = t2_cm; W.nrec[slot][1] = t2_cl; W.nrec[slot][2] = t2_lm;
                    }
                    st_pair += __popc(t2_need);
                    __syncwarp();
                    for(int i0 = 0; i0 < nf; i0 += S) {
                        const int idx = i0 + s;
                        const bool act = idx < nf;
                        const int ii = act ? idx : nf - 1;
                        const float4 a = W.nrec[ii][0], c4 = W.nrec[ii][1], lm = W.nrec[ii][2];
                        const float dr2 = a.w, de = c4.w, len = lm.x;
                        const bool mine = act && ((__float_as_uint(lm.z) >> t) & 1u);
                        const float dx = a.x - tx, dy = a.y - ty, dz = a.z - tz;
                        const float r2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                        bool k_lo = true, k_hi = true;
                        // M l^2 > r^4 a  <=>  r^2 < sqrt(M l^2) / sqrt(a): one threshold per (node, target), its rounding
                        // (sqrt, rsqrt.approx, products: < 1e-6) inside the relative slack, the coordinate error added
                        const float th = lm.y * rta;
                        bool o_lo = r2 < fmaf(th, 1.f + SQ_EPS, dr2), o_hi = r2 < fmaf(th, 1.f - SQ_EPS, -dr2);
                        if(lm.w == 0.f) {
                            const float eff = fmaf(0.5f, len, P.f_rcut), ins = 0.6f * len, bhq = len * len * P.f_invbh2;
                            const float rcA = P.f_rc2_hi + dr2, rcB = P.f_rc2_lo - dr2;
                            const float efA = fmaf(eff, 1.f + SQ_EPS, de), efB = fmaf(eff, 1.f - SQ_EPS, -de);
                            const float inA = fmaf(ins, 1.f + SQ_EPS, de), inB = fmaf(ins, 1.f - SQ_EPS, -de);
                            const float bhA = fmaf(bhq, 1.f + SQ_EPS, dr2), bhB = fmaf(bhq, 1.f - SQ_EPS, -dr2);
                            const float em = fmaxf(fmaxf(fabsf(c4.x - tx), fabsf(c4.y - ty)), fabsf(c4.z - tz));
                            k_lo = (r2 <= rcA) || !(em > efA);
                            k_hi = (r2 <= rcB) || !(em > efB);
                            o_lo = o_lo || (r2 < bhA) || (em < inA);
                            o_hi = o_hi || (r2 < bhB) || (em < inB);
                        }
                        const bool ex = k_lo != k_hi || (k_lo && o_lo != o_hi);
                        const unsigned bx = __ballot_sync(SQ_FULL, mine && ex);
                        const unsigned bo = __ballot_sync(SQ_FULL, mine && !ex && k_lo && o_lo);
                        const unsigned ba = __ballot_sync(SQ_FULL, mine && !ex && k_lo && !o_lo);
                        if(act && t == 0) {
                            const unsigned GM = G == 32 ? 0xffffffffu : ((1u << (G & 31)) - 1u);
                            W.nrec[idx][2] = make_float4(__uint_as_float((bo >> (G * s)) & GM), __uint_as_float((ba >> (G * s)) & GM),
                                                         __uint_as_float((bx >> (G * s)) & GM), 0.f);
                        }
                    }
                    __syncwarp();
                    if(t2_need) {
                        const float4 r = W.nrec[slot][2];
                        open_mask |= __float_as_uint(r.x); acc_mask |= __float_as_uint(r.y); risky_exact |= __float_as_uint(r.z);
                    }
                    __syncwarp();
                }
            }
            // ---- tier 3: the reference's FP64 expressions on the FP64 node and particle data
            if(risky_exact) {
                unsigned exact = risky_exact;
                const double4 M = cm[node], C = cl[node];
                sq_exact_par X;
                X.box = P.box; X.half = P.half; X.rcut = P.rcut; X.rcut2 = P.rcut2; X.bh2 = P.bh2; X.relative = P.relative;
                while(exact) {
                    const int t2 = 31 - __clz(exact);
                    const unsigned bit = 1u << t2;
                    exact ^= bit;
                    const long long ti = targets ? (long long) targets[t0 + t2] : (t0 + t2);
                    const int c = sq_exact_class(M, C, tpos[3 * ti], tpos[3 * ti + 1], tpos[3 * ti + 2], P.errtol * toldacc[ti], X);
                    st_exact++;
                    if(c == 1) acc_mask |= bit;
                    else if(c == 2) open_mask |= bit;
                }
            }
            const int type = SQ_INFO_TYPE(L.w);
            // a pseudo node some target opens: the reference exports that particle (gravshort2.hpp:400-411); here the
            // sub-tree should have been imported below it - counted so that the caller can tell
            if(open_mask && type == 2) st_pseudo += __popc(open_mask);
            // ---- opened internal nodes -> push children with the open mask (pseudo nodes are skipped: gravshort2.hpp:305-314)
            {
                const unsigned cmk = SQ_INFO_CHILDMASK(L.w);
                unsigned sv = cmk;
                if(open_mask && type == 1 && intkeep) {
                    // children whose keep box (Rcut + len/2 around the cell centre, gravshort2.hpp:160) misses the whole
                    // group are dropped here, before anyone loads them: the reference would discard them for every
                    // target.  Child centres = this centre +- len/4, in quanta to within 2; e_child carries the slack.
                    const int lvl = SQ_INFO_LEVEL(L.w);
                    const int ec = P.e_child[lvl], q = P.q4[lvl];
                    if(ec >= 0) {
                        const int c0 = (int) (nodeC.x - ix0), c1 = (int) (nodeC.y - iy0), c2 = (int) (nodeC.z - iz0);
                        const int x0 = sq_sub_sat(c0, q), x1 = sq_add_sat(c0, q), y0 = sq_sub_sat(c1, q), y1 = sq_add_sat(c1, q), z0 = sq_sub_sat(c2, q), z1 = sq_add_sat(c2, q);
                        unsigned okx = 0, oky = 0, okz = 0;
                        if(!(sq_sub_sat(x0, ec) > mxx || sq_add_sat(x0, ec) < mnx)) okx |= 0x55u;
                        if(!(sq_sub_sat(x1, ec) > mxx || sq_add_sat(x1, ec) < mnx)) okx |= 0xAAu;
                        if(!(sq_sub_sat(y0, ec) > mxy || sq_add_sat(y0, ec) < mny)) oky |= 0x33u;
                        if(!(sq_sub_sat(y1, ec) > mxy || sq_add_sat(y1, ec) < mny)) oky |= 0xCCu;
                        if(!(sq_sub_sat(z0, ec) > mxz || sq_add_sat(z0, ec) < mnz)) okz |= 0x0Fu;
                        if(!(sq_sub_sat(z1, ec) > mxz || sq_add_sat(z1, ec) < mnz)) okz |= 0xF0u;
                        sv &= okx & oky & okz;
                    }
                }
                const bool pushes = open_mask && type == 1 && sv != 0u;
                const unsigned bp = __ballot_sync(SQ_FULL, pushes);
                if(pushes) {
                    const int p = top + __popc(bp & sq_lanemask_lt());
                    W.stack[p] = make_int2(L.x, (int) open_mask);
                    W.scnt[p] = (unsigned char) sv;
                    W.scm[p] = (unsigned char) cmk;
                }
                top += __popc(bp);
            }
            // ---- opened leaves -> their particles are sources (gravshort2.hpp:291-303)
            int tot_leaf;
            {
                const int np = (open_mask && type == 0) ? L.y : 0;
                const int off = sq_warp_excl_scan(np, tot_leaf);
                if(tot_leaf) {
                    st_int += (unsigned long long) np * __popc(open_mask);
                    W.s_first[lane] = L.x - off;        // particle = s_first + j
                    W.s_mask[lane] = open_mask;
                    for(int k = 0; k < np; k++) W.src[off + k] = (unsigned char) lane;
                }
            }
            st_int += __popc(acc_mask);
            const unsigned anyacc = __ballot_sync(SQ_FULL, acc_mask != 0u);
            __syncwarp();
            // ---- sources join the lists, 32 per round: round 0 = accepted nodes, then the leaf particles
            const int nrounds = 1 + ((tot_leaf + 31) >> 5);
            // leaf particles are fetched one round ahead of their use (the load latency hides behind the previous round)
            uint4 qn = make_uint4(0u, 0u, 0u, 0u);
            unsigned mn = 0u;
#if SQ_PREFETCH
            if((int) lane < tot_leaf) {
                const int lo = W.src[lane];
                qn = ipart[(long long) W.s_first[lo] + (int) lane];
                mn = W.s_mask[lo];
            }
#endif
            for(int rd = 0; rd < nrounds; rd++) {
                bool v = false;
                float4 e = nodeE;
                unsigned m = acc_mask;
                if(rd == 0) {
                    if(!drain && anyacc == 0u) continue;
                    v = acc_mask != 0u;
                } else {
#if SQ_PREFETCH
                    const uint4 q4 = qn;
                    m = mn;
                    v = ((rd - 1) << 5) + (int) lane < tot_leaf;
                    const int jn = (rd << 5) + (int) lane;
                    if(jn < tot_leaf) {
                        const int lo = W.src[jn];
                        qn = ipart[(long long) W.s_first[lo] + jn];
                        mn = W.s_mask[lo];
                    }
#else
                    const int j = ((rd - 1) << 5) + (int) lane;
                    v = j < tot_leaf;
                    uint4 q4 = qn;
                    if(v) {
                        const int lo = W.src[j];
                        q4 = ipart[(long long) W.s_first[lo] + j];
                        m = W.s_mask[lo];
                    }
#endif
                    if(v) e = make_float4((float) (int) (q4.x - gcix) * P.qs, (float) (int) (q4.y - gciy) * P.qs, (float) (int) (q4.z - gciz) * P.qs, __uint_as_float(q4.w));
                }
                // ---- near list
                const bool near = v && (pairwrap || (fabsf(e.x) <= nhx && fabsf(e.y) <= nhy && fabsf(e.z) <= nhz));
                const unsigned bn = __ballot_sync(SQ_FULL, near);
                if(bn != 0u || (drain && nn > 0)) {
                    const int cn = __popc(bn);
                    if(nn + cn > NEAR_CAP || drain) {
                        __syncwarp();
                        sq_eval_near<POT, S>(W, nn, tabrow0, P, pairwrap, tx, ty, tz, s, t, ax, ay, az, ap);
                        st_ent += (nn + S - 1) / S;
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
                // ---- interaction list
                const bool isfar = v && !near;
                const unsigned bf = __ballot_sync(SQ_FULL, isfar);
                const int cf = __popc(bf);
                if(nlist + cf > LIST_CAP || (drain && nlist > 0)) {
                    // pad to whole groups of 4 steps: far away, reaching nobody
                    const int nsteps4 = ((nlist + S - 1) / S + 3) & ~3;
                    for(int kk = nlist + (int) lane; kk < nsteps4 * S; kk += 32) {
                        const int mp = sq_list_slot<S>(kk);
                        W.lx[mp] = SQ_FAR; W.ly[mp] = SQ_FAR; W.lz[mp] = SQ_FAR; W.lm[mp] = 0.f;
                        sq_store_field<G>(W.mw, kk, 0u);
                    }
                    __syncwarp();
#ifndef SQ_NOEVAL
                    sq_eval_list<POT, S>(W, nsteps4, tabaddr, tx, ty, tz, s, lanebit, ax, ay, az, ap);
#endif
                    st_ent += nsteps4;
                    dax += (double) ax; day += (double) ay; daz += (double) az;
                    ax = ay = az = 0.f;
                    if(POT) { dap += (double) ap; ap = 0.f; }
                    nlist = 0;
                    __syncwarp();
                }
                if(isfar) {
                    const int p = nlist + __popc(bf & sq_lanemask_lt());
                    const int mp = sq_list_slot<S>(p);
                    W.lx[mp] = e.x; W.ly[mp] = e.y; W.lz[mp] = e.z; W.lm[mp] = e.w;
                    sq_store_field<G>(W.mw, p, m);
                }
                nlist += cf;
            }
            __syncwarp();
            if(drain) break;
        }
        dax += (double) ax; day += (double) ay; daz += (double) az;
        if(POT) dap += (double) ap;
        // combine the slices of each target
        #pragma unroll
        for(int o = G; o < 32; o <<= 1) {
            dax += __shfl_xor_sync(SQ_FULL, dax, o);
            day += __shfl_xor_sync(SQ_FULL, day, o);
            daz += __shfl_xor_sync(SQ_FULL, daz, o);
            if(POT) dap += __shfl_xor_sync(SQ_FULL, dap, o);
        }
        if(gcost) {
            // interactions of this group: the work measure the domain decomposition balances (domain.cpp:452-527 uses GravCost)
            const unsigned gi = __reduce_add_sync(SQ_FULL, (unsigned) (st_int - st_int_before));
            if(lane == 0) gcost[g] = gi;
        }
        if(valid && s == 0) {
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
        st_pseudo += __shfl_xor_sync(SQ_FULL, st_pseudo, o);
    }
    if(lane == 0) {
        atomicAdd(&stats[0], st_int);
        atomicAdd(&stats[1], st_ent);
        atomicAdd(&stats[2], st_tst);
        atomicAdd(&stats[3], st_pair);
        atomicAdd(&stats[4], st_exact);
        if(st_pseudo) atomicAdd(&stats[5], st_pseudo);
    }
}

// ---- per-walk preparation: fixed-point node and source records
__global__ void sq_walk_prep_nodes_kernel(const double4 * __restrict__ cm, const double4 * __restrict__ cl, long long n,
                                          uint4 * __restrict__ icm, uint4 * __restrict__ icl, double qinv, double s, double minv)
{
    for(long long i = blockIdx.x * (long long) blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const double4 M = cm[i], C = cl[i];
        icm[i] = make_uint4(sq_quant(M.x, qinv), sq_quant(M.y, qinv), sq_quant(M.z, qinv), __float_as_uint((float) (M.w * minv)));
        icl[i] = make_uint4(sq_quant(C.x, qinv), sq_quant(C.y, qinv), sq_quant(C.z, qinv), __float_as_uint((float) (C.w * s)));
    }
}
__global__ void sq_walk_prep_parts_kernel(const double * __restrict__ pos, const float * __restrict__ mass, long long n,
                                          uint4 * __restrict__ ipart, double qinv, double minv)
{
    for(long long i = blockIdx.x * (long long) blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x)
        ipart[i] = make_uint4(sq_quant(pos[3 * i], qinv), sq_quant(pos[3 * i + 1], qinv), sq_quant(pos[3 * i + 2], qinv),
                              __float_as_uint((float) ((double) mass[i] * minv)));
}
// ------------------------------------------------------------------------------------ host side
static const double h_sr_x[SQ_NGRAVTAB] = SQ_SHORTRANGE_X_INIT;
static const double h_sr_pot[SQ_NGRAVTAB] = SQ_SHORTRANGE_POT_INIT;
static const double h_sr_force[SQ_NGRAVTAB] = SQ_SHORTRANGE_FORCE_INIT;

// GravShortTable, libgadget/gravshort-tree2.cpp:29-69: float tables; lookup gravity.h:50-60 (i = floor(r / dx),
// linear interpolation between rows i and i+1, nothing from row 511 on).  Two device forms, TAB_ROWS float2 rows:
//   force only:  row i = (intercept at r = 0, slope) of the line through points i and i+1, rows 511.. = (0, 0)
//   force + pot: row i = (force intercept, force slope, potential intercept, potential slope), rows 511.. = 0
static int sq_walk_table(sq_ctx * c)
{
    // the device image only changes with (window type, Asmth) or a new raw table: keep it across walks (hierarchical
    // sub-steps call the walk thousands of times per PM step)
    const uint64_t sig = c->have_raw ? (0x8000000000000000ull | c->raw_serial) : (((uint64_t) c->par.WindowType << 56) ^ (uint64_t) llround(c->par.Asmth * 1e9));
    if(c->w_table.p && c->w_table_sig == sig) return SQ_OK;
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
    // device image: TAB_ROWS float2 rows (force only) followed by TAB_ROWS float4 rows (force + potential)
    static_assert(TAB_ROWS % 2 == 1, "float4 part starts 8 bytes after a 16-byte boundary otherwise");
    struct { float2 f[TAB_ROWS + 1]; float4 fp[TAB_ROWS]; } h;
    memset(&h, 0, sizeof(h));
    for(int i = 0; i + 1 < SQ_NTAB; i++) {
        const double sf = (double) ft[i + 1] - (double) ft[i], sp = (double) pt[i + 1] - (double) pt[i];
        // the line through rows i and i+1 as (intercept at r = 0, slope): value = intercept + r * slope for r in [i, i+1)
        const float fsf = (float) sf, fsp = (float) sp;
        h.f[i] = make_float2((float) ((double) ft[i] - (double) i * (double) fsf), fsf);
        h.fp[i] = make_float4((float) ((double) ft[i] - (double) i * (double) fsf), fsf, (float) ((double) pt[i] - (double) i * (double) fsp), fsp);
    }
    SQ_TRY(c->w_table.reserve(sizeof(h)));
    // pageable source: the call returns once `h` has been staged, no stream synchronisation needed
    SQ_CUDA(cudaMemcpyAsync(c->w_table.p, &h, sizeof(h), cudaMemcpyHostToDevice, c->stream));
    c->w_table_sig = sig;
    return SQ_OK;
}

template <bool POT, int G>
static int sq_walk_launch(sq_ctx * c, const uint32_t * targets, int64_t ntargets, const sq_walk_par & P)
{
    const size_t smem = TAB_BYTES + WALK_WARPS * sizeof(sq_walk_warp_smem);
    SQ_CUDA(cudaFuncSetAttribute(sq_walk_kernel<POT, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    int per_sm = 0;
    SQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sq_walk_kernel<POT, G>, WALK_THREADS, smem));
    if(per_sm < 1) per_sm = 1;
    const int64_t ngroups = (ntargets + G - 1) / G;
    int64_t grid = (int64_t) c->num_sms * per_sm;
    const int64_t need = (ngroups + WALK_WARPS - 1) / WALK_WARPS;
    if(grid > need) grid = need;
    if(grid < 1) grid = 1;
    sq_walk_kernel<POT, G><<<(unsigned) grid, WALK_THREADS, smem, c->stream>>>(c->w_icm.as<uint4>(), c->w_icl.as<uint4>(), c->t_lk.as<int4>(),
            c->t_cm.as<double4>(), c->t_cl.as<double4>(), c->w_ipart.as<uint4>(), c->pos.as<double>(), c->mass.as<float>(), c->oldacc.as<double>(),
            targets, (long long) ntargets, c->w_table.as<float2>() + (POT ? TAB_ROWS + 1 : 0), c->w_accel.as<double>(), c->w_pot.as<double>(),
            c->w_stats.as<unsigned long long>(), c->w_counter.as<int>(), c->w_gcost.as<unsigned>(), P);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

int sq_walk_group_width = 32;   // tunable: targets per warp (32, 16 or 8)

// cost[upload index] = interactions of the particle's group / group width, for the targets of the last walk
__global__ void sq_walk_cost_kernel(const unsigned * __restrict__ gcost, const uint32_t * __restrict__ targets, const uint32_t * __restrict__ perm,
                                    long long m, int G, float * __restrict__ out)
{
    for(long long q = blockIdx.x * (long long) blockDim.x + threadIdx.x; q < m; q += (long long) gridDim.x * blockDim.x) {
        const long long i = targets ? targets[q] : q;
        out[perm[i]] = (float) gcost[q / G] / (float) G;
    }
}

extern "C" int sq_walk_particle_cost(sq_ctx * c, float * cost)
{
    if(!c || !cost || !c->walk_done) { sq_set_error("sq_walk_particle_cost: no walk results"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaSetDevice(c->device));
    const int64_t n = c->n, m = c->walk_ntargets;
    sq_buf tmp;
    SQ_TRY(tmp.reserve(sizeof(float) * (size_t) (n + 1)));
    cudaError_t e = cudaMemsetAsync(tmp.p, 0, sizeof(float) * (size_t) (n + 1), c->stream);
    if(e == cudaSuccess && m > 0) {
        sq_walk_cost_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(c->w_gcost.as<unsigned>(), c->walk_all ? nullptr : c->w_targets.as<uint32_t>(), c->perm.as<uint32_t>(),
                (long long) m, c->walk_group, tmp.as<float>());
        c->kernel_launches++;
        e = cudaGetLastError();
    }
    if(e == cudaSuccess) e = cudaMemcpyAsync(cost, tmp.p, sizeof(float) * (size_t) n, cudaMemcpyDefault, c->stream);
    if(e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    tmp.release();
    if(e != cudaSuccess) { sq_set_error("sq_walk_particle_cost: %s", cudaGetErrorString(e)); return SQ_ERR_CUDA; }
    return SQ_OK;
}

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
    const double munit = c->walk_munit > 0 ? c->walk_munit : ((c->tree_total_mass > 0 && c->tree_np > 0) ? c->tree_total_mass / (double) c->tree_np : 1.0);
    P.minv = 1.0 / munit;
    P.Ka = P.G * munit * P.s * P.s;
    P.Kp = munit * P.s;
    const int G = sq_walk_group_width;
    if(G != 8 && G != 16 && G != 32) { sq_set_error("sq_tree_walk: group width must be 8, 16 or 32"); return SQ_ERR_ARG; }
    P.qinv = 4294967296.0 / P.box;
    P.qs = (float) (P.box * P.s / 4294967296.0);
    P.f_rcut = (float) (P.rcut * P.s);
    P.f_rc2_lo = (float) (P.rcut2 * P.s * P.s) * (1.f - SQ_EPS);
    P.f_rc2_hi = (float) (P.rcut2 * P.s * P.s) * (1.f + SQ_EPS);
    P.f_invbh2 = (float) (1.0 / P.bh2);
    P.f_hs = (float) (P.h * P.s);
    P.f_h2 = P.f_hs * P.f_hs;
    P.f_boxs = (float) (P.box * P.s);
    P.f_invboxs = (float) (1.0 / (P.box * P.s));
    P.f_halfs = (float) (P.half * P.s);
    P.f_range = (float) (SQ_NTAB - 1);
    // keep box per level in fixed-point quanta: eff = Rcut + len/2 (gravshort2.hpp:160), len = 1.001 Box / 2^level
    for(int L = 0; L < SQ_MAXLEVEL + 3; L++) {
        const double len = ldexp(1.001 * par.BoxSize, -L);
        const double effq = (P.rcut + 0.5 * len) * P.qinv;
        if(effq >= 2147483648.0 + 8.0) { P.e_in[L] = -1; P.e_out[L] = -1; }              // >= Box/2: |NEAREST()| never exceeds it
        else if(effq > 2147483648.0 - 8.0 || effq < 8.0) { P.e_in[L] = -2; P.e_out[L] = -2; }
        else { P.e_in[L] = (int) floor(effq) - 2; P.e_out[L] = (int) ceil(effq) + 2; }
    }
    for(int L = 0; L < SQ_MAXLEVEL + 3; L++) {
        const double q = 0.25 * ldexp(1.001 * par.BoxSize, -L) * P.qinv;
        const bool ok = L + 1 < SQ_MAXLEVEL + 3 && P.e_out[L + 1] >= 0 && P.e_out[L + 1] < (1 << 30) - 16 && q < (double) (1 << 30);
        P.q4[L] = ok ? (int) floor(q) : 0;
        P.e_child[L] = ok ? P.e_out[L + 1] + 8 : -1;
    }

    const int64_t n = c->n;
    const uint32_t * targets = nullptr;
    int64_t ntargets = c->n_valid;
    c->walk_all = !active && c->range_count < 0;
    if(active) {
        SQ_TRY(sq_active_to_sorted(c, active, nactive, &c->w_targets, &ntargets));
        targets = c->w_targets.as<uint32_t>();
    } else if(c->range_count >= 0) {
        // sq_tree_walk_range: this rank's share of a globally sorted set
        SQ_TRY(sq_walk_range_targets(c, &c->w_targets));
        targets = c->w_targets.as<uint32_t>();
        ntargets = c->range_count;
    }
    const bool pot = c->tree_full;     // update_potential = tree->full_particle_tree_flag (gravshort-tree2.cpp:148)
    SQ_TRY(c->w_accel.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
    SQ_TRY(c->w_pot.reserve(sizeof(double) * (size_t) (n + 1)));
    SQ_TRY(c->w_stats.reserve(64));
    SQ_TRY(c->w_counter.reserve(64));
    SQ_TRY(c->w_gcost.reserve(sizeof(unsigned) * (size_t) (ntargets / 8 + 2)));
    c->walk_group = G;
    SQ_CUDA(cudaMemsetAsync(c->w_stats.p, 0, 64, c->stream));
    SQ_CUDA(cudaMemsetAsync(c->w_counter.p, 0, 64, c->stream));
    SQ_CUDA(cudaMemsetAsync(c->w_accel.p, 0, sizeof(double) * 3 * (size_t) n, c->stream));
    if(pot) SQ_CUDA(cudaMemsetAsync(c->w_pot.p, 0, sizeof(double) * (size_t) n, c->stream));
    SQ_CUDA(cudaEventRecord(c->ev[8], c->stream));
    if(ntargets > 0) {
        // fixed-point node and source records for this walk (scales depend on the call's parameters)
        const double * spos = c->tree_identity ? c->pos.as<double>() : c->tree_pos.as<double>();
        const float * smass = c->tree_identity ? c->mass.as<float>() : c->tree_mass.as<float>();
        SQ_TRY(c->w_icm.reserve(sizeof(uint4) * (size_t) (c->tree_numnodes + 1)));
        SQ_TRY(c->w_icl.reserve(sizeof(uint4) * (size_t) (c->tree_numnodes + 1)));
        const int64_t nsrc = c->let_attached ? c->n_sources : c->tree_np;      // local + imported particles
        SQ_TRY(c->w_ipart.reserve(sizeof(uint4) * (size_t) (nsrc + 1)));
        sq_walk_prep_nodes_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(c->t_cm.as<double4>(), c->t_cl.as<double4>(), (long long) c->tree_numnodes,
                c->w_icm.as<uint4>(), c->w_icl.as<uint4>(), P.qinv, P.s, P.minv);
        sq_walk_prep_parts_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(spos, smass, (long long) nsrc, c->w_ipart.as<uint4>(), P.qinv, P.minv);
        c->kernel_launches += 2;
        SQ_CUDA(cudaGetLastError());
        int rc;
        if(pot) rc = G == 8 ? sq_walk_launch<true, 8>(c, targets, ntargets, P) : (G == 16 ? sq_walk_launch<true, 16>(c, targets, ntargets, P) : sq_walk_launch<true, 32>(c, targets, ntargets, P));
        else rc = G == 8 ? sq_walk_launch<false, 8>(c, targets, ntargets, P) : (G == 16 ? sq_walk_launch<false, 16>(c, targets, ntargets, P) : sq_walk_launch<false, 32>(c, targets, ntargets, P));
        SQ_TRY(rc);
    }
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
        stats->ngroups = (ntargets + G - 1) / G;
        stats->interactions = (int64_t) h_stats[0];
        stats->list_entries = (int64_t) h_stats[1];
        stats->nodes_tested = (int64_t) h_stats[2];
        stats->pair_tests = (int64_t) h_stats[3];
        stats->exact_tests = (int64_t) h_stats[4];
        stats->pseudo_opened = (int64_t) h_stats[5];
        stats->kernel_ms = ms;
    }
    // TreeUseBH > 1: Barnes-Hut on the first call only (gravshort-tree2.cpp:170-171)
    if(c->par.TreeUseBH > 1 && !c->have_raw) c->par.TreeUseBH = 0;
    return SQ_OK;
}
