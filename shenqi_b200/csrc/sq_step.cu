// the steps either side of each tree call in hierarchical stepping, on device arrays, so that
// particles stay resident between sub-steps (BASELINE config 5).  All three are HBM-bound streaming kernels.
//
//   sq_drift_device           real_drift_particle / drift_all_particles    libgadget/drift.cpp:16-78, 81-99
//   sq_kick_device            apply_hierarchical_grav_kick + do_grav_short_range_kick
//                                                                          libgadget/timestep.cpp:247-288, 962-968
//   sq_active_sublist_device  SubActivePredicate + build_active_sublist    libgadget/timestep.cpp:1353-1397
//                             is_timebin_active                            libgadget/timestep.cpp:132-139
//
// The reference is compiled for x86-64 without FMA contraction (Options.mk.example:9): products and sums are rounded
// separately here as well (__dmul_rn / __dadd_rn), so positions and velocities agree bit for bit.  The drift sums in the
// order (Pos + shift) + Vel*ddrift, which is what the reference's -ffast-math build makes of drift.cpp:69 (checked
// against oracle/_ref and tests/golden/step_ref.npz); the source order differs from it by at most 1 ulp.
#include "sq_common.cuh"

#define SQS_TILE 2048        // items per block in the compaction: 8 warps x 8 rounds x 32 lanes

static int step_grid(const sq_ctx * c, long long n, int per_sm)
{
    long long need = (n + 255) / 256, cap = (long long) c->num_sms * per_sm;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}

// One thread per coordinate (fully coalesced over the [n][3] arrays); algorithmic bytes: 48 read + 24 written per particle.
// drift.cpp:20-29: garbage / swallowed particles only follow the random shift.  drift.cpp:68-75: a non-finite position is
// fatal in the reference (endrun 5): counted in *bad, reported by the host wrapper.  drift.cpp:76-79: periodic wrap into (0, Box].
__global__ void __launch_bounds__(256) sqs_drift_kernel(double * __restrict__ pos, const double * __restrict__ vel, const unsigned char * __restrict__ skip,
                                                        long long n3, double ddrift, double sx, double sy, double sz, double box, unsigned long long * bad)
{
    const long long stride = (long long) gridDim.x * blockDim.x;
    for(long long e = (long long) blockIdx.x * blockDim.x + threadIdx.x; e < n3; e += stride) {
        const long long i = e / 3;
        const int j = (int) (e - 3 * i);
        const double sh = j == 0 ? sx : (j == 1 ? sy : sz);
        double p = pos[e];
        if(skip && skip[i])
            p = __dadd_rn(p, sh);
        else {
            p = __dadd_rn(__dadd_rn(p, sh), __dmul_rn(vel[e], ddrift));
            if(!isfinite(p)) { atomicAdd(bad, 1ull); continue; }
        }
        while(p > box) p = __dadd_rn(p, -box);
        while(p <= 0) p = __dadd_rn(p, box);
        pos[e] = p;
    }
}

// timestep.cpp:272-279, 962-968: Vel[pa] += Accel[pa] * gravkick for every active pa that is neither swallowed nor garbage.
// One thread per (active particle, axis); 4 B index + 48 B read + 24 B written per active particle.
__global__ void __launch_bounds__(256) sqs_kick_kernel(double * __restrict__ vel, const double * __restrict__ accel, const int * __restrict__ active,
                                                       const unsigned char * __restrict__ skip, long long na3, double gravkick)
{
    const long long stride = (long long) gridDim.x * blockDim.x;
    for(long long e = (long long) blockIdx.x * blockDim.x + threadIdx.x; e < na3; e += stride) {
        const long long i = e / 3;
        const int j = (int) (e - 3 * i);
        const long long pa = active ? (long long) active[i] : i;
        if(skip && skip[pa]) continue;
        const long long q = 3 * pa + j;
        vel[q] = __dadd_rn(vel[q], __dmul_rn(accel[q], gravkick));
    }
}

// timestep.cpp:1359-1369 with is_timebin_active (timestep.cpp:132-139) and dti_from_timebin (timebinmgr.h:42-45).
__device__ __forceinline__ bool sqs_sub_active(int pi, const unsigned char * __restrict__ timebin, const unsigned char * __restrict__ skip,
                                               int maxtimebin, long long ti_current)
{
    if(skip && skip[pi]) return false;
    const int bin = timebin[pi];
    if(bin > maxtimebin) return false;
    if(bin <= 0 || ti_current <= 0) return true;
    return (ti_current & ((1ll << bin) - 1)) == 0;      // current % 2^bin == 0 for current > 0
}

// A warp owns 256 consecutive candidates as 8 rounds of 32 lanes; the order inside the warp is (round, lane), so
// output order = input order (std::copy_if is stable, timestep.cpp:1382-1390).
template <bool SCATTER>
__global__ void __launch_bounds__(256) sqs_sublist_kernel(const int * __restrict__ active, long long nact, const unsigned char * __restrict__ timebin,
                                                          const unsigned char * __restrict__ skip, int maxtimebin, long long ti_current,
                                                          int * __restrict__ block_counts, int * __restrict__ out)
{
    __shared__ int wsum[8];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const long long base = (long long) blockIdx.x * SQS_TILE + w * 256;
    unsigned bal[8];
    int pis[8];
    int total = 0;
    #pragma unroll
    for(int r = 0; r < 8; r++) {
        const long long i = base + r * 32 + lane;
        bool keep = false;
        int pi = 0;
        if(i < nact) {
            pi = active ? active[i] : (int) i;
            keep = sqs_sub_active(pi, timebin, skip, maxtimebin, ti_current);
        }
        pis[r] = pi;
        bal[r] = __ballot_sync(SQ_FULL, keep);
        total += __popc(bal[r]);
    }
    if(lane == 0) wsum[w] = total;
    __syncthreads();
    if(!SCATTER) {
        if(threadIdx.x == 0) {
            int s = 0;
            #pragma unroll
            for(int k = 0; k < 8; k++) s += wsum[k];
            block_counts[blockIdx.x] = s;
        }
        return;
    }
    int off = block_counts[blockIdx.x];      // exclusive offsets after sqs_scan_kernel
    for(int k = 0; k < w; k++) off += wsum[k];
    #pragma unroll
    for(int r = 0; r < 8; r++) {
        if((bal[r] >> lane) & 1u) out[off + __popc(bal[r] & sq_lanemask_lt())] = pis[r];
        off += __popc(bal[r]);
    }
}

// Exclusive scan of the block counts in place by one block (n / 2048 entries: 8192 for 256^3), total to *total.
__global__ void __launch_bounds__(1024) sqs_scan_kernel(int * __restrict__ counts, int nblocks, long long * total)
{
    __shared__ int wtot[32];
    __shared__ int carry_s;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if(threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for(int b0 = 0; b0 < nblocks; b0 += 1024) {
        const int i = b0 + threadIdx.x;
        const int v = i < nblocks ? counts[i] : 0;
        int wt;
        int x = sq_warp_excl_scan(v, wt);
        if(lane == 0) wtot[w] = wt;
        __syncthreads();
        const int carry = carry_s;
        int woff = 0, all = 0;
        for(int k = 0; k < 32; k++) { if(k < w) woff += wtot[k]; all += wtot[k]; }
        if(i < nblocks) counts[i] = carry + woff + x;
        __syncthreads();
        if(threadIdx.x == 0) carry_s = carry + all;
        __syncthreads();
    }
    if(threadIdx.x == 0) *total = carry_s;
}

// timestep.cpp:351-377 per active particle.  Every FP64 operation is rounded separately in source order (IEEE division and
// square root), so the integer step and the bin equal the CPU restatement's.
__device__ __forceinline__ long long sqs_gravity_dti(const sq_timebin_params & P, const double * __restrict__ ga, const double * __restrict__ pm, double * dloga_out)
{
    // grav_acceleration2, timestep.cpp:1012-1027
    const double a2inv = __ddiv_rn(1.0, __dmul_rn(P.atime, P.atime));
    const double ax = __dadd_rn(__dmul_rn(a2inv, ga[0]), __dmul_rn(a2inv, pm[0]));
    const double ay = __dadd_rn(__dmul_rn(a2inv, ga[1]), __dmul_rn(a2inv, pm[1]));
    const double az = __dadd_rn(__dmul_rn(a2inv, ga[2]), __dmul_rn(a2inv, pm[2]));
    double ac2 = __dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az));
    if(ac2 == 0) ac2 = 1.0e-60;
    // get_timestep_gravity_dloga, timestep.cpp:1029-1039
    const double ac = __dsqrt_rn(ac2);
    const double dt = __dsqrt_rn(__ddiv_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, P.ErrTolIntAccuracy), P.atime), P.softening), ac));
    double dloga = __dmul_rn(dt, P.hubble);
    *dloga_out = dloga;
    // convert_timestep_to_ti, timestep.cpp:157-174
    if(P.dti_max == 0) return 0;
    if(dloga < P.MinSizeTimestep) dloga = P.MinSizeTimestep;
    // TimeBinMgr::dti_from_dloga + ti_from_loga_snap, timebinmgr.h:123-156
    const double target = __dadd_rn(dloga, P.loga);
    long long snap = P.lastsnap;
    int k = 0;
    if(P.can_advance && P.sync_loga[1] <= target) { snap++; k = 1; }
    const double logDTime = __ddiv_rn(__dadd_rn(P.sync_loga[k + 1], -P.sync_loga[k]), (double) (1ull << SQ_TIMEBINS));
    long long tip = snap << SQ_TIMEBINS;
    const double tnew = __dadd_rn((double) tip, __ddiv_rn(__dadd_rn(target, -P.sync_loga[k]), logDTime));
    // `ti += double` converts back to inttime_t; the x86 conversion of an out-of-range or NaN value gives INT64_MIN
    tip = (tnew >= 9223372036854775808.0 || tnew < -9223372036854775808.0 || tnew != tnew) ? (long long) 0x8000000000000000ull : (long long) tnew;
    long long dti = tip - P.Ti_Current;
    if(dti > P.dti_max || dti < 0) dti = P.dti_max;
    return dti;
}

__global__ void __launch_bounds__(256) sqs_timebin_kernel(sq_timebin_params P, const double * __restrict__ accel, const double * __restrict__ gravpm,
                                                          const int * __restrict__ active, const unsigned char * __restrict__ skip, long long nactive,
                                                          unsigned char * __restrict__ timebin, unsigned long long * __restrict__ counts /* [48]: bins, then nbad */)
{
    __shared__ unsigned int hist[SQ_TIMEBINS + 2];
    for(int k = threadIdx.x; k < SQ_TIMEBINS + 2; k += blockDim.x) hist[k] = 0;
    __syncthreads();
    const long long stride = (long long) gridDim.x * blockDim.x;
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nactive; i += stride) {
        const long long pa = active ? (long long) active[i] : i;
        if(skip && skip[pa]) continue;
        double dloga;
        long long dti = sqs_gravity_dti(P, accel + 3 * pa, gravpm + 3 * pa, &dloga);
        // round_down_power_of_two (timebinmgr.cpp:190) and get_timestep_bin (timestep.cpp:1236-1250)
        if(dti > 0) dti = 1ll << (63 - __clzll(dti));      // largest power of two <= dti ...
        if(dti > (1ll << SQ_TIMEBINS)) dti = 1ll << SQ_TIMEBINS;   // ... starting from TIMEBASE (timebinmgr.cpp:193-200)
        if(dti <= 1 || dti > (1ll << SQ_TIMEBINS)) atomicAdd(&hist[SQ_TIMEBINS + 1], 1u);
        int bin = dti <= 1 ? 0 : 63 - __clzll(dti);
        if(bin > P.largest_active) bin = P.largest_active;
        timebin[pa] = (unsigned char) bin;
        atomicAdd(&hist[bin], 1u);
    }
    __syncthreads();
    for(int k = threadIdx.x; k < SQ_TIMEBINS + 2; k += blockDim.x)
        if(hist[k]) atomicAdd(&counts[k], (unsigned long long) hist[k]);
}

__global__ void __launch_bounds__(256) sqs_pushdown_kernel(const int * __restrict__ active, long long nactive, unsigned char * __restrict__ timebin, int push_down_bin)
{
    const long long stride = (long long) gridDim.x * blockDim.x;
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < nactive; i += stride) {
        const long long pa = active ? (long long) active[i] : i;
        if(timebin[pa] > push_down_bin) timebin[pa] = (unsigned char) push_down_bin;
    }
}

extern "C" {

int sq_drift_device(sq_ctx * c, double * pos, const double * vel, const unsigned char * skip, int64_t n, double ddrift,
                    const double * random_shift, double BoxSize)
{
    if(!c || n < 0 || (n > 0 && (!pos || !vel)) || !(BoxSize > 0)) { sq_set_error("sq_drift_device: bad argument"); return SQ_ERR_ARG; }
    if(n == 0) return SQ_OK;
    SQ_CUDA(cudaSetDevice(c->device));
    SQ_TRY(c->step_tmp.reserve(64));
    unsigned long long * bad = c->step_tmp.as<unsigned long long>();
    SQ_CUDA(cudaMemsetAsync(bad, 0, sizeof(*bad), c->stream));
    const double s0 = random_shift ? random_shift[0] : 0, s1 = random_shift ? random_shift[1] : 0, s2 = random_shift ? random_shift[2] : 0;
    sqs_drift_kernel<<<step_grid(c, 3 * n, 16), 256, 0, c->stream>>>(pos, vel, skip, 3 * (long long) n, ddrift, s0, s1, s2, BoxSize, bad);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    unsigned long long hbad = 0;
    SQ_CUDA(cudaMemcpyAsync(&hbad, bad, sizeof(hbad), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    if(hbad) { sq_set_error("sq_drift_device: %llu non-finite coordinates after the drift (drift.cpp:70-73 is fatal)", hbad); return SQ_ERR_NONFINITE; }
    return SQ_OK;
}

int sq_kick_device(sq_ctx * c, double * vel, const double * accel, const int32_t * active, const unsigned char * skip, int64_t nactive, double gravkick)
{
    if(!c || nactive < 0 || (nactive > 0 && (!vel || !accel))) { sq_set_error("sq_kick_device: bad argument"); return SQ_ERR_ARG; }
    if(nactive == 0) return SQ_OK;
    SQ_CUDA(cudaSetDevice(c->device));
    sqs_kick_kernel<<<step_grid(c, 3 * nactive, 16), 256, 0, c->stream>>>(vel, accel, active, skip, 3 * (long long) nactive, gravkick);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

int sq_active_sublist_device(sq_ctx * c, const int32_t * active, int64_t nactive, const unsigned char * timebin, const unsigned char * skip,
                             int maxtimebin, int64_t Ti_Current, int32_t * sublist, int64_t * nsub)
{
    if(!c || !nsub || nactive < 0 || nactive > 0x7fffffffLL || (nactive > 0 && (!timebin || !sublist))) {
        sq_set_error("sq_active_sublist_device: bad argument");
        return SQ_ERR_ARG;
    }
    *nsub = 0;
    if(nactive == 0) return SQ_OK;
    SQ_CUDA(cudaSetDevice(c->device));
    const int nblocks = sq_div_up(nactive, SQS_TILE);
    SQ_TRY(c->step_tmp.reserve(64 + sizeof(int) * (size_t) nblocks));
    long long * total = c->step_tmp.as<long long>() + 1;
    int * counts = (int *) (c->step_tmp.as<char>() + 64);
    sqs_sublist_kernel<false><<<nblocks, 256, 0, c->stream>>>(active, nactive, timebin, skip, maxtimebin, Ti_Current, counts, nullptr);
    sqs_scan_kernel<<<1, 1024, 0, c->stream>>>(counts, nblocks, total);
    sqs_sublist_kernel<true><<<nblocks, 256, 0, c->stream>>>(active, nactive, timebin, skip, maxtimebin, Ti_Current, counts, sublist);
    c->kernel_launches += 3;
    SQ_CUDA(cudaGetLastError());
    long long h = 0;
    SQ_CUDA(cudaMemcpyAsync(&h, total, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    *nsub = h;
    return SQ_OK;
}

int sq_gravity_timebins_device(sq_ctx * c, const sq_timebin_params * par, const double * accel, const double * gravpm, const int32_t * active,
                               const unsigned char * skip, int64_t nactive, unsigned char * timebin, int64_t * counts, int64_t * nbad)
{
    if(!c || !par || !counts || nactive < 0 || (nactive > 0 && (!accel || !gravpm || !timebin)) || par->largest_active < 0 ||
       par->largest_active > SQ_TIMEBINS || !(par->atime > 0)) {
        sq_set_error("sq_gravity_timebins_device: bad argument");
        return SQ_ERR_ARG;
    }
    for(int k = 0; k <= SQ_TIMEBINS; k++) counts[k] = 0;
    if(nbad) *nbad = 0;
    if(nactive == 0) return SQ_OK;
    SQ_CUDA(cudaSetDevice(c->device));
    SQ_TRY(c->step_tmp.reserve(64 + 8 * (SQ_TIMEBINS + 2)));
    unsigned long long * dcounts = (unsigned long long *) (c->step_tmp.as<char>() + 64);
    SQ_CUDA(cudaMemsetAsync(dcounts, 0, 8 * (SQ_TIMEBINS + 2), c->stream));
    sqs_timebin_kernel<<<step_grid(c, nactive, 8), 256, 0, c->stream>>>(*par, accel, gravpm, active, skip, nactive, timebin, dcounts);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    unsigned long long h[SQ_TIMEBINS + 2];
    SQ_CUDA(cudaMemcpyAsync(h, dcounts, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    for(int k = 0; k <= SQ_TIMEBINS; k++) counts[k] = (int64_t) h[k];
    if(nbad) *nbad = (int64_t) h[SQ_TIMEBINS + 1];
    return SQ_OK;
}

int sq_timebin_pushdown_device(sq_ctx * c, const int32_t * active, int64_t nactive, unsigned char * timebin, int push_down_bin)
{
    if(!c || nactive < 0 || (nactive > 0 && !timebin) || push_down_bin < 0 || push_down_bin > SQ_TIMEBINS) { sq_set_error("sq_timebin_pushdown_device: bad argument"); return SQ_ERR_ARG; }
    if(nactive == 0) return SQ_OK;
    SQ_CUDA(cudaSetDevice(c->device));
    sqs_pushdown_kernel<<<step_grid(c, nactive, 8), 256, 0, c->stream>>>(active, nactive, timebin, push_down_bin);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

} // extern "C"
