// Device-resident entry points of the multi-GPU path (one rank per GPU).
//
// The exchanges themselves (NCCL all-gather / all-to-all / send-recv over NVLink) are issued by the host layer
// (shenqi_b200/dist.py, torch.distributed); this file holds what runs on each GPU between them and takes plain
// device pointers, so the same C ABI serves a C++ host with raw NCCL:
//   * tree side: install a globally PH-sorted particle set that arrived by all-gather (sq_set_sorted_device), walk
//     only this rank's range of it (sq_tree_walk_range).  Stands where the reference merges the top tree and
//     exports/imports walks (forcetree.cpp:1144-1282, treewalk2.h:282-373, 742-812): here every rank holds the whole
//     (fixed-point) tree.  TEST MODE since round 2: memory and traffic are O(N_total) per GPU (512^3 fits, 1024^3 does
//     not); the production tree path is the domain-decomposed one of sq_let.cu (O(N/P + boundary)).
//   * PM side: x-slab decomposition of the mesh.  Particles are routed to the owner of their x-plane
//     (sq_pm_slab_route; the reference moves painted cells instead, petapm.cpp:692-1107), painted into the slab plus
//     one ghost plane (sq_pm_slab_paint; petapm.cpp:1132-1183, 1304-1310), transformed as 2-D r2c per plane ->
//     all-to-all transpose -> 1-D along x (sq_pm_slab_fft_*; petapm_fft_r2c/c2r petapm.cpp:49-71 through heFFTe),
//     multiplied by the gravity transfer functions (sq_pm_slab_transfer; gravpm.cpp:378-487) and read out with the
//     four fields fused (sq_pm_slab_readout; gravpm.cpp:489-500).
#include "sq_common.cuh"
#include <math.h>

int sq_iota(sq_ctx * c, uint32_t * v, int64_t n);
static int dist_grid(const sq_ctx * c, long long n, int per_sm)
{
    long long need = (n + 255) / 256, cap = (long long) c->num_sms * per_sm;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}

__device__ __forceinline__ int sqd_wrap(int i, int n) { return i >= n ? i - n : (i < 0 ? i + n : i); }
// petapm.cpp:1146-1158: t = Pos/CellSize, iCell = floor(t), Res = t - iCell
__device__ __forceinline__ void sqd_cic_cell(double p, double cell, int nmesh, int & i0, double & res)
{
    const double t = p / cell;
    const double f = floor(t);
    res = t - f;
    int i = (int) f;
    i %= nmesh;
    if(i < 0) i += nmesh;
    i0 = i;
}

// ------------------------------------------------------------------------------------------------ routing
__global__ void __launch_bounds__(256) sqd_route_count_kernel(const double * __restrict__ pos, long long n, double cell, int nmesh, int nxl,
                                                               int * __restrict__ counts, unsigned char * __restrict__ owner)
{
    __shared__ int h[64];
    if(threadIdx.x < 64) h[threadIdx.x] = 0;
    __syncthreads();
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        int ix; double r;
        sqd_cic_cell(pos[3 * i], cell, nmesh, ix, r);
        const int o = ix / nxl;
        owner[i] = (unsigned char) o;
        atomicAdd(&h[o], 1);
    }
    __syncthreads();
    if(threadIdx.x < 64 && h[threadIdx.x]) atomicAdd(&counts[threadIdx.x], h[threadIdx.x]);
}

// records (x, y, z, m) grouped by owner; slot[i] = where particle i went (results come back to the same slot)
__global__ void __launch_bounds__(256) sqd_route_scatter_kernel(const double * __restrict__ pos, const float * __restrict__ mass, long long n,
                                                                 const unsigned char * __restrict__ owner, int * __restrict__ cursor,
                                                                 double4 * __restrict__ send, int * __restrict__ slot)
{
    // A CTA takes tiles of 2048 consecutive particles: one global reservation per owner per tile (the cursors are few
    // and hot), the warps then claim their runs from the CTA's share in shared memory.  Particles that were
    // neighbours stay within one tile-sized run of the output, so paint and readout keep their cache locality.
    constexpr int TILE = 2048;
    __shared__ int h[64], cur[64];
    const unsigned lane = sq_lane();
    const int warp = threadIdx.x >> 5;
    const long long ntiles = (n + TILE - 1) / TILE;
    for(long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long t0 = tile * TILE;
        if(threadIdx.x < 64) h[threadIdx.x] = 0;
        __syncthreads();
        for(int k = threadIdx.x; k < TILE; k += 256)
            if(t0 + k < n) atomicAdd(&h[owner[t0 + k]], 1);
        __syncthreads();
        if(threadIdx.x < 64) cur[threadIdx.x] = h[threadIdx.x] ? atomicAdd(&cursor[threadIdx.x], h[threadIdx.x]) : 0;
        __syncthreads();
        for(int k0 = warp * 32; k0 < TILE; k0 += 256) {
            const long long i = t0 + k0 + lane;
            const bool ok = i < n;
            const int o = ok ? (int) owner[i] : -1 - (int) lane;
            const unsigned peers = sq_match64((long long) o, (unsigned) o & 255u);
            const int leader = __ffs(peers) - 1;
            int base = 0;
            if(ok && (int) lane == leader) base = atomicAdd(&cur[o], __popc(peers));
            base = __shfl_sync(SQ_FULL, base, leader);
            if(ok) {
                const int p = base + __popc(peers & sq_lanemask_lt());
                send[p] = make_double4(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], (double) mass[i]);
                slot[i] = p;
            }
        }
        __syncthreads();
    }
}

// out[i] = in[slot[i]] (4 doubles each): results of the routed particles back in the caller's order
__global__ void sqd_unroute_kernel(const double4 * __restrict__ in, const int * __restrict__ slot, long long n, double4 * __restrict__ out)
{
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x)
        out[i] = in[slot[i]];
}

// ------------------------------------------------------------------------------------------------ slab CIC
// slab = (nxl + 1) x N x N doubles, z fastest: planes x0 .. x0 + nxl - 1 of the mesh, then the ghost plane x0 + nxl
// (owned by the next rank; added there after the paint, fetched from there before the readout).
__global__ void __launch_bounds__(256) sqd_slab_paint_kernel(const double4 * __restrict__ part, long long n, double cell, int nmesh, int x0,
                                                             double * __restrict__ slab)
{
    const unsigned lane = sq_lane();
    const long long nit = (n + 31) / 32;
    const long long w0 = ((long long) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long) gridDim.x * blockDim.x) >> 5;
    for(long long wi = w0; wi < nit; wi += nw) {
        const long long i = wi * 32 + lane;
        const bool ok = i < n;
        int ix = 0, iy = 0, iz = 0;
        double rx = 0, ry = 0, rz = 0, m = 0;
        if(ok) {
            const double4 p = part[i];
            sqd_cic_cell(p.x, cell, nmesh, ix, rx);
            sqd_cic_cell(p.y, cell, nmesh, iy, ry);
            sqd_cic_cell(p.z, cell, nmesh, iz, rz);
            m = p.w;
        }
        const int lx = ix - x0;
        const int iy1 = sqd_wrap(iy + 1, nmesh), iz1 = sqd_wrap(iz + 1, nmesh);
        #pragma unroll
        for(int c = 0; c < 8; c++) {
            const int ox = c & 1, oy = (c >> 1) & 1, oz = (c >> 2) & 1;
            // petapm.cpp:1165-1176: weight = 1.0 * wx * wy * wz
            double w = 1.0;
            w *= ox ? rx : (1 - rx);
            w *= oy ? ry : (1 - ry);
            w *= oz ? rz : (1 - rz);
            const double v = w * m;
            const long long lin = ok ? ((long long) (lx + ox) * nmesh + (oy ? iy1 : iy)) * nmesh + (oz ? iz1 : iz) : -1 - (long long) lane;
            // warp-aggregated: lanes that hit the same cell are summed, one red.add.f64 per distinct cell
            const unsigned peers = sq_match64(lin, (unsigned) ((lx + ox) & 7) | ((unsigned) ((oy ? iy1 : iy) & 7) << 3) | ((unsigned) ((oz ? iz1 : iz) & 3) << 6));
            unsigned rem = peers;
            double sum = 0;
            while(__any_sync(SQ_FULL, rem != 0)) {
                const int src = rem ? (__ffs(rem) - 1) : (int) lane;
                const double got = __shfl_sync(SQ_FULL, v, src);
                if(rem) { sum += got; rem &= rem - 1; }
            }
            if(ok && (peers & sq_lanemask_lt()) == 0)
                atomicAdd(&slab[lin], sum);
        }
    }
}

// out[i] = {potential, fx, fy, fz} at particle i from the four slabs (ghost planes filled)
__global__ void __launch_bounds__(256) sqd_slab_readout_kernel(const double4 * __restrict__ part, long long n, double cell, int nmesh, int x0,
                                                               const double * __restrict__ m0, const double * __restrict__ m1,
                                                               const double * __restrict__ m2, const double * __restrict__ m3,
                                                               double4 * __restrict__ out)
{
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const double4 p = part[i];
        int ix, iy, iz;
        double rx, ry, rz;
        sqd_cic_cell(p.x, cell, nmesh, ix, rx);
        sqd_cic_cell(p.y, cell, nmesh, iy, ry);
        sqd_cic_cell(p.z, cell, nmesh, iz, rz);
        const int lx = ix - x0;
        const int iy1 = sqd_wrap(iy + 1, nmesh), iz1 = sqd_wrap(iz + 1, nmesh);
        double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        #pragma unroll
        for(int c = 0; c < 8; c++) {
            const int ox = c & 1, oy = (c >> 1) & 1, oz = (c >> 2) & 1;
            double w = 1.0;
            w *= ox ? rx : (1 - rx);
            w *= oy ? ry : (1 - ry);
            w *= oz ? rz : (1 - rz);
            const long long lin = ((long long) (lx + ox) * nmesh + (oy ? iy1 : iy)) * nmesh + (oz ? iz1 : iz);
            a0 += w * m0[lin];
            a1 += w * m1[lin];
            a2 += w * m2[lin];
            a3 += w * m3[lin];
        }
        out[i] = make_double4(a0, a1, a2, a3);
    }
}

// ------------------------------------------------------------------------------------------------ transposes
// forward: planes [nxl][N][Nzh] -> blocks for the all-to-all [P][nxl][nyl][Nzh] (block q = the y-range of rank q)
__global__ void sqd_pack_kernel(const double2 * __restrict__ in, int nxl, int nmesh, int nyl, int nzh, double2 * __restrict__ out)
{
    const long long total = (long long) nxl * nmesh * nzh;
    for(long long ip = (long long) blockIdx.x * blockDim.x + threadIdx.x; ip < total; ip += (long long) gridDim.x * blockDim.x) {
        const int z = (int) (ip % nzh);
        const long long t = ip / nzh;
        const int y = (int) (t % nmesh), x = (int) (t / nmesh);
        const int q = y / nyl, yl = y - q * nyl;
        out[(((long long) q * nxl + x) * nyl + yl) * nzh + z] = in[ip];
    }
}
// inverse: received blocks [P][nxl][nyl][Nzh] (block q = y-range of rank q) -> planes [nxl][N][Nzh]
__global__ void sqd_unpack_kernel(const double2 * __restrict__ in, int nxl, int nmesh, int nyl, int nzh, double2 * __restrict__ out)
{
    const long long total = (long long) nxl * nmesh * nzh;
    for(long long ip = (long long) blockIdx.x * blockDim.x + threadIdx.x; ip < total; ip += (long long) gridDim.x * blockDim.x) {
        const int z = (int) (ip % nzh);
        const long long t = ip / nzh;
        const int y = (int) (t % nmesh), x = (int) (t / nmesh);
        const int q = y / nyl, yl = y - q * nyl;
        out[ip] = in[(((long long) q * nxl + x) * nyl + yl) * nzh + z];
    }
}

// One pass over this rank's modes [kx][ky in y0..y0+nyl)[kz]: phi_k from rho_k (gravpm.cpp:378-434), then the three
// force components (gravpm.cpp:448-487).  isinc2[k] = 1/sinc^2(pi k/N), dker[k] = -diff_kernel(2 pi k/N) * N/Box.
__global__ void __launch_bounds__(256) sqd_slab_transfer_kernel(const double2 * rho, int nmesh, int y0, int nyl, double asmth2, double pot_factor,
                                                                const double * __restrict__ isinc2, const double * __restrict__ dker,
                                                                double2 * potk, double2 * __restrict__ fxk,
                                                                double2 * __restrict__ fyk, double2 * __restrict__ fzk)
{
    const int nz = nmesh / 2 + 1;
    const long long total = (long long) nmesh * nyl * nz;
    for(long long ip = (long long) blockIdx.x * blockDim.x + threadIdx.x; ip < total; ip += (long long) gridDim.x * blockDim.x) {
        const int iz = (int) (ip % nz);
        const long long t = ip / nz;
        const int iy = y0 + (int) (t % nyl);
        const int ix = (int) (t / nyl);
        // petapm.cpp:159-162
        const int kx = ix <= nmesh / 2 ? ix : ix - nmesh;
        const int ky = iy <= nmesh / 2 ? iy : iy - nmesh;
        const int kz = iz;
        const long long k2 = (long long) kx * kx + (long long) ky * ky + (long long) kz * kz;
        double2 v = rho[ip];
        if(k2 == 0) {
            v.x = 0; v.y = 0;
        } else {
            const double smth = exp(-(double) k2 * asmth2) / (double) k2;
            double f = 1.0;
            f *= isinc2[ix];
            f *= isinc2[iy];
            f *= isinc2[iz];
            const double fac = pot_factor * smth * f * f;
            v.x *= fac; v.y *= fac;
        }
        potk[ip] = v;
        if(fxk) {
            const double fx = dker[ix], fy = dker[iy], fz = dker[iz];
            fxk[ip] = make_double2(-v.y * fx, v.x * fx);
            fyk[ip] = make_double2(-v.y * fy, v.x * fy);
            fzk[ip] = make_double2(-v.y * fz, v.x * fz);
        }
    }
}

__global__ void sqd_range_kernel(uint32_t * v, long long n, uint32_t first)
{
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x)
        v[i] = first + (uint32_t) i;
}

// ------------------------------------------------------------------------------------------------ host side
static double h_sinc_unnormed_d(double x)
{
    if(x < 1e-5 && x > -1e-5) {
        const double x2 = x * x;
        return 1.0 - x2 / 6. + x2 * x2 / 120.;
    }
    return sin(x) / x;
}

void sq_dist_destroy(sq_ctx * c)
{
    if(c->slab_r2c) cufftDestroy(c->slab_r2c);
    if(c->slab_c2r) cufftDestroy(c->slab_c2r);
    if(c->slab_x) cufftDestroy(c->slab_x);
    c->slab_r2c = c->slab_c2r = c->slab_x = 0;
    c->slab_N = c->slab_P = 0;
}

static int slab_check(sq_ctx * c, int nranks, const char * who)
{
    if(!c || !c->have_params) { sq_set_error("%s: call sq_set_params first", who); return SQ_ERR_ARG; }
    const int N = c->par.Nmesh;
    if(nranks < 1 || nranks > 64 || N % nranks != 0 || (N & 1)) {
        sq_set_error("%s: Nmesh %d must be even and divisible by the number of ranks %d (<= 64)", who, N, nranks);
        return SQ_ERR_ARG;
    }
    return SQ_OK;
}

static int slab_plans(sq_ctx * c, int nranks)
{
    const int N = c->par.Nmesh;
    if(c->slab_N == N && c->slab_P == nranks) return SQ_OK;
    sq_dist_destroy(c);
    const int nxl = N / nranks, nyl = N / nranks, nzh = N / 2 + 1;
    int n2[2] = {N, N};
    SQ_CUFFT(cufftPlanMany(&c->slab_r2c, 2, n2, nullptr, 1, 0, nullptr, 1, 0, CUFFT_D2Z, nxl));
    SQ_CUFFT(cufftPlanMany(&c->slab_c2r, 2, n2, nullptr, 1, 0, nullptr, 1, 0, CUFFT_Z2D, nxl));
    // along x of [N][nyl][nzh]: element stride nyl*nzh, one transform per (y, z)
    int n1[1] = {N};
    const int str = nyl * nzh;
    SQ_CUFFT(cufftPlanMany(&c->slab_x, 1, n1, n1, str, 1, n1, str, 1, CUFFT_Z2Z, str));
    SQ_CUFFT(cufftSetStream(c->slab_r2c, c->stream));
    SQ_CUFFT(cufftSetStream(c->slab_c2r, c->stream));
    SQ_CUFFT(cufftSetStream(c->slab_x, c->stream));
    // per-axis factors: gravpm.cpp:395-401 (CIC deconvolution) and :436-460 (differencing)
    std::vector<double> tab(2 * (size_t) N);
    for(int i = 0; i < N; i++) {
        const int k = i <= N / 2 ? i : i - N;
        double tmp = h_sinc_unnormed_d((k * M_PI) / N);
        tab[i] = 1. / (tmp * tmp);
        const double w = k * (2 * M_PI / N);
        tab[N + i] = -1 * (1 / 6.0 * (8 * sin(w) - sin(2 * w))) * (N / c->par.BoxSize);
    }
    SQ_TRY(c->slab_tab.reserve(sizeof(double) * 2 * (size_t) N));
    SQ_CUDA(cudaMemcpyAsync(c->slab_tab.p, tab.data(), sizeof(double) * 2 * (size_t) N, cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    c->slab_N = N; c->slab_P = nranks;
    return SQ_OK;
}

extern "C" {

int sq_ctx_set_stream(sq_ctx * c, void * stream)
{
    if(!c) return SQ_ERR_ARG;
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    if(c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    c->stream = (cudaStream_t) stream;
    c->own_stream = false;
    if(c->plan_r2c) { SQ_CUFFT(cufftSetStream(c->plan_r2c, c->stream)); SQ_CUFFT(cufftSetStream(c->plan_c2r, c->stream)); }
    if(c->slab_r2c) { SQ_CUFFT(cufftSetStream(c->slab_r2c, c->stream)); SQ_CUFFT(cufftSetStream(c->slab_c2r, c->stream)); SQ_CUFFT(cufftSetStream(c->slab_x, c->stream)); }
    return SQ_OK;
}

int sq_device_ptr(sq_ctx * c, int which, void ** ptr, int64_t * count)
{
    if(!c || !ptr) return SQ_ERR_ARG;
    void * p = nullptr;
    int64_t n = c->n;
    switch(which) {
    case SQ_PTR_SORTED_KEYS: p = c->sorted ? c->keys.p : nullptr; break;
    case SQ_PTR_SORTED_POS: p = c->sorted ? c->pos.p : nullptr; break;
    case SQ_PTR_SORTED_MASS: p = c->sorted ? c->mass.p : nullptr; break;
    case SQ_PTR_SORTED_PERM: p = c->sorted ? c->perm.p : nullptr; break;
    case SQ_PTR_WALK_ACCEL: p = c->walk_done ? c->w_accel.p : nullptr; break;
    case SQ_PTR_WALK_POT: p = (c->walk_done && c->tree_full) ? c->w_pot.p : nullptr; break;
    case SQ_PTR_PM_OUT: p = c->pm_done ? c->pm_out.p : nullptr; break;
    default: sq_set_error("sq_device_ptr: unknown array %d", which); return SQ_ERR_ARG;
    }
    if(!p) { sq_set_error("sq_device_ptr: array %d is not available yet (call order)", which); return SQ_ERR_ARG; }
    *ptr = p;
    if(count) *count = n;
    return SQ_OK;
}

int sq_upload_particles_device(sq_ctx * c, const double * pos, const float * mass, int64_t n)
{
    if(!c || (n > 0 && (!pos || !mass))) { sq_set_error("sq_upload_particles_device: null argument"); return SQ_ERR_ARG; }
    if(n < 0 || n >= ((int64_t) 1 << 31)) { sq_set_error("particle count %lld out of range (MaxPart < 2^31, partmanager.c:22)", (long long) n); return SQ_ERR_ARG; }
    c->n = n;
    c->n_valid = n; c->have_skip = false;
    c->sorted = false; c->tree_built = false; c->walk_done = false; c->pm_done = false;
    SQ_TRY(c->pos_in.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
    SQ_TRY(c->mass_in.reserve(sizeof(float) * (size_t) (n + 1)));
    SQ_TRY(c->oldacc_in.reserve(sizeof(double) * (size_t) (n + 1)));
    SQ_CUDA(cudaMemsetAsync(c->oldacc_in.p, 0, sizeof(double) * (size_t) (n + 1), c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->pos_in.p, pos, sizeof(double) * 3 * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->mass_in.p, mass, sizeof(float) * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
    return SQ_OK;
}

int sq_get_geo_mismatch(sq_ctx * c, int * flag)
{
    if(!c || !flag) return SQ_ERR_ARG;
    *flag = c->geo_mismatch;
    return SQ_OK;
}

// A PH-sorted particle set that is already on the device (the all-gather of every rank's sorted range, in rank order
// = global key order when the ranks own consecutive key ranges, as after the reference's domain decomposition).
int sq_set_sorted_device(sq_ctx * c, const uint64_t * keys, const double * pos, const float * mass, int64_t n, int geo_mismatch)
{
    if(!c || !c->have_params) { sq_set_error("sq_set_sorted_device: call sq_set_params first"); return SQ_ERR_ARG; }
    if(n < 0 || n >= ((int64_t) 1 << 31)) { sq_set_error("particle count %lld out of range (MaxPart < 2^31, partmanager.c:22)", (long long) n); return SQ_ERR_ARG; }
    if(n > 0 && (!keys || !pos || !mass)) { sq_set_error("sq_set_sorted_device: null argument"); return SQ_ERR_ARG; }
    c->n = n;
    c->n_valid = n; c->have_skip = false;
    c->sorted = false; c->tree_built = false; c->walk_done = false; c->pm_done = false;
    SQ_TRY(c->keys.reserve(sizeof(uint64_t) * (size_t) (n + 1)));
    SQ_TRY(c->pos.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
    SQ_TRY(c->mass.reserve(sizeof(float) * (size_t) (n + 1)));
    SQ_TRY(c->oldacc.reserve(sizeof(double) * (size_t) (n + 1)));
    SQ_TRY(c->perm.reserve(sizeof(uint32_t) * (size_t) (n + 1)));
    SQ_TRY(c->rank.reserve(sizeof(uint32_t) * (size_t) (n + 1)));
    SQ_TRY(c->t_counters.reserve(256));
    SQ_CUDA(cudaMemsetAsync(c->t_counters.p, 0, 256, c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->keys.p, keys, sizeof(uint64_t) * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->pos.p, pos, sizeof(double) * 3 * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->mass.p, mass, sizeof(float) * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaMemsetAsync(c->oldacc.p, 0, sizeof(double) * (size_t) (n + 1), c->stream));
    SQ_TRY(sq_iota(c, c->perm.as<uint32_t>(), n));
    SQ_TRY(sq_iota(c, c->rank.as<uint32_t>(), n));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    c->geo_mismatch = geo_mismatch ? 1 : 0;
    c->sorted = true;
    return SQ_OK;
}

// |FullTreeGravAccel + GravPM| / G of the sorted particles first .. first + count - 1 (this rank's targets)
int sq_set_oldacc_range_device(sq_ctx * c, int64_t first, int64_t count, const double * oldacc)
{
    if(!c || !c->sorted || first < 0 || count < 0 || first + count > c->n || (count > 0 && !oldacc)) { sq_set_error("sq_set_oldacc_range_device: bad range"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaMemcpyAsync(c->oldacc.as<double>() + first, oldacc, sizeof(double) * (size_t) count, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    return SQ_OK;
}

// grav_short_tree for the sorted particles first .. first + count - 1 only; results stay on the device
// (SQ_PTR_WALK_ACCEL / SQ_PTR_WALK_POT, indexed by sorted position).
int sq_tree_walk_range(sq_ctx * c, int64_t first, int64_t count, sq_walk_stats * stats)
{
    if(!c || !c->sorted || first < 0 || count < 0 || first + count > c->n) { sq_set_error("sq_tree_walk_range: bad range"); return SQ_ERR_ARG; }
    c->range_first = first; c->range_count = count;
    const int rc = sq_walk_run(c, nullptr, 0, stats);
    c->range_count = -1;
    return rc;
}

} /* extern "C" */

int sq_walk_range_targets(sq_ctx * c, sq_buf * out)
{
    SQ_TRY(out->reserve(sizeof(uint32_t) * (size_t) (c->range_count + 1)));
    if(c->range_count > 0) {
        sqd_range_kernel<<<dist_grid(c, c->range_count, 8), 256, 0, c->stream>>>(out->as<uint32_t>(), c->range_count, (uint32_t) c->range_first);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    return SQ_OK;
}

extern "C" {

// Group this rank's particles by the owner of their x-plane (rank = ix / (Nmesh / nranks)).
// send4: n x 4 doubles (x, y, z, m) grouped by destination rank; slot[i] = index of particle i in send4;
// counts[r] (host) = particles going to rank r.
int sq_pm_slab_route(sq_ctx * c, const double * pos, const float * mass, int64_t n, int nranks, double * send4, int32_t * slot, int64_t * counts)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_route"));
    if(n > 0 && (!pos || !mass || !send4 || !slot)) { sq_set_error("sq_pm_slab_route: null argument"); return SQ_ERR_ARG; }
    const int N = c->par.Nmesh, nxl = N / nranks;
    const double cell = c->par.BoxSize / N;
    SQ_TRY(c->route_tmp.reserve((size_t) n + 1024));
    int * d_counts = c->route_tmp.as<int>();                          // 64 counts + 64 cursors
    unsigned char * owner = c->route_tmp.as<unsigned char>() + 512;
    SQ_CUDA(cudaMemsetAsync(d_counts, 0, 512, c->stream));
    int h[64];
    memset(h, 0, sizeof(h));
    if(n > 0) {
        sqd_route_count_kernel<<<dist_grid(c, n, 8), 256, 0, c->stream>>>(pos, n, cell, N, nxl, d_counts, owner);
        c->kernel_launches++;
        SQ_CUDA(cudaMemcpyAsync(h, d_counts, sizeof(int) * 64, cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        int cur[64], acc = 0;
        for(int r = 0; r < 64; r++) { cur[r] = acc; acc += h[r]; }
        SQ_CUDA(cudaMemcpyAsync(d_counts + 64, cur, sizeof(int) * 64, cudaMemcpyHostToDevice, c->stream));
        sqd_route_scatter_kernel<<<dist_grid(c, n, 8), 256, 0, c->stream>>>(pos, mass, n, owner, d_counts + 64, (double4 *) send4, slot);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
        SQ_CUDA(cudaStreamSynchronize(c->stream));     // cur is a host temporary
    }
    for(int r = 0; r < nranks; r++) counts[r] = h[r];
    return SQ_OK;
}

// out4[i] = back4[slot[i]]: the routed particles' results in the caller's particle order
int sq_pm_slab_unroute(sq_ctx * c, const double * back4, const int32_t * slot, int64_t n, double * out4)
{
    if(!c || (n > 0 && (!back4 || !slot || !out4))) return SQ_ERR_ARG;
    if(n > 0) {
        sqd_unroute_kernel<<<dist_grid(c, n, 8), 256, 0, c->stream>>>((const double4 *) back4, slot, n, (double4 *) out4);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    return SQ_OK;
}

// Zero the slab ((Nmesh/nranks + 1) x Nmesh x Nmesh doubles) and paint the routed particles (part4: m x 4 doubles).
int sq_pm_slab_paint(sq_ctx * c, const double * part4, int64_t m, int rank, int nranks, double * slab)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_paint"));
    const int N = c->par.Nmesh, nxl = N / nranks;
    SQ_CUDA(cudaEventRecord(c->ev[10], c->stream));
    SQ_CUDA(cudaMemsetAsync(slab, 0, sizeof(double) * (size_t) (nxl + 1) * N * N, c->stream));
    if(m > 0) {
        sqd_slab_paint_kernel<<<dist_grid(c, m, 8), 256, 0, c->stream>>>((const double4 *) part4, m, c->par.BoxSize / N, N, rank * nxl, slab);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    SQ_CUDA(cudaEventRecord(c->ev[11], c->stream));
    return SQ_OK;
}

// 2-D r2c of the slab's nxl planes (work: nxl x N x Nzh complex) and packing for the transpose
// (packed: nranks blocks of nxl x nyl x Nzh complex, block q goes to rank q).
int sq_pm_slab_fft_fwd(sq_ctx * c, double * slab, int nranks, double * work, double * packed)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_fft_fwd"));
    SQ_TRY(slab_plans(c, nranks));
    const int N = c->par.Nmesh, nxl = N / nranks, nzh = N / 2 + 1;
    SQ_CUFFT(cufftExecD2Z(c->slab_r2c, slab, (cufftDoubleComplex *) work));
    sqd_pack_kernel<<<dist_grid(c, (long long) nxl * N * nzh, 16), 256, 0, c->stream>>>((const double2 *) work, nxl, N, N / nranks, nzh, (double2 *) packed);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

// 1-D transform along x of data[N][nyl][Nzh] in place; inverse != 0: unnormalised inverse (heFFTe scale::none)
int sq_pm_slab_fft_x(sq_ctx * c, double * data, int nranks, int inverse)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_fft_x"));
    SQ_TRY(slab_plans(c, nranks));
    SQ_CUFFT(cufftExecZ2Z(c->slab_x, (cufftDoubleComplex *) data, (cufftDoubleComplex *) data, inverse ? CUFFT_INVERSE : CUFFT_FORWARD));
    return SQ_OK;
}

// rho_k[N][nyl][Nzh] (ky = rank * nyl ..) -> out4: four arrays of the same shape, back to back: phi_k, Fx_k, Fy_k, Fz_k
int sq_pm_slab_transfer(sq_ctx * c, const double * rho, int rank, int nranks, double * out4)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_transfer"));
    SQ_TRY(slab_plans(c, nranks));
    const int N = c->par.Nmesh, nyl = N / nranks, nzh = N / 2 + 1;
    const size_t per = (size_t) N * nyl * nzh;
    const double asmth2 = pow((2 * M_PI) * c->par.Asmth / N, 2);
    const double pot_factor = -c->par.G / (M_PI * c->par.BoxSize);
    double2 * o = (double2 *) out4;
    sqd_slab_transfer_kernel<<<dist_grid(c, (long long) per, 16), 256, 0, c->stream>>>((const double2 *) rho, N, rank * nyl, nyl, asmth2, pot_factor,
            c->slab_tab.as<double>(), c->slab_tab.as<double>() + N, o, o + per, o + 2 * per, o + 3 * per);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

// rho_k -> phi_k in place of the four outputs: the forces are differenced in real space (sq_pm_slab_gradient)
int sq_pm_slab_transfer_phi(sq_ctx * c, const double * rho, int rank, int nranks, double * phik)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_transfer_phi"));
    SQ_TRY(slab_plans(c, nranks));
    const int N = c->par.Nmesh, nyl = N / nranks, nzh = N / 2 + 1;
    const size_t per = (size_t) N * nyl * nzh;
    const double asmth2 = pow((2 * M_PI) * c->par.Asmth / N, 2);
    const double pot_factor = -c->par.G / (M_PI * c->par.BoxSize);
    sqd_slab_transfer_kernel<<<dist_grid(c, (long long) per, 16), 256, 0, c->stream>>>((const double2 *) rho, N, rank * nyl, nyl, asmth2, pot_factor,
            c->slab_tab.as<double>(), c->slab_tab.as<double>() + N, (double2 *) phik, nullptr, nullptr, nullptr);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

// Potential slab (planes 0 .. nxl-1 valid) + the neighbours' planes -> force slabs by the 4-point difference.
//   lo2: planes -2, -1 (the previous rank's last two), hi3: planes nxl .. nxl+2 (the next rank's first three)
//   ext: work space (nxl + 5) planes; fx, fy, fz: (nxl + 1) planes each (the last one is the CIC ghost plane);
//   phi[nxl] is filled with the ghost plane as well.
int sq_pm_slab_gradient(sq_ctx * c, double * phi, const double * lo2, const double * hi3, int nranks, double * ext, double * fx, double * fy, double * fz)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_gradient"));
    const int N = c->par.Nmesh, nxl = N / nranks;
    if(nxl < 3) { sq_set_error("sq_pm_slab_gradient: needs at least 3 planes per rank"); return SQ_ERR_ARG; }
    const size_t plane = (size_t) N * N;
    SQ_CUDA(cudaMemcpyAsync(ext, lo2, sizeof(double) * 2 * plane, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(ext + 2 * plane, phi, sizeof(double) * (size_t) nxl * plane, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(ext + (size_t) (2 + nxl) * plane, hi3, sizeof(double) * 3 * plane, cudaMemcpyDeviceToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(phi + (size_t) nxl * plane, hi3, sizeof(double) * plane, cudaMemcpyDeviceToDevice, c->stream));
    return sq_pm_gradient(c, ext + 2 * plane, nxl + 1, 0, fx, fy, fz);
}

// received blocks [nranks][nxl][nyl][Nzh] -> planes (work) -> 2-D c2r into the slab's nxl planes
int sq_pm_slab_fft_inv(sq_ctx * c, const double * recv, int nranks, double * work, double * slab)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_fft_inv"));
    SQ_TRY(slab_plans(c, nranks));
    const int N = c->par.Nmesh, nxl = N / nranks, nzh = N / 2 + 1;
    sqd_unpack_kernel<<<dist_grid(c, (long long) nxl * N * nzh, 16), 256, 0, c->stream>>>((const double2 *) recv, nxl, N, N / nranks, nzh, (double2 *) work);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    SQ_CUFFT(cufftExecZ2D(c->slab_c2r, (cufftDoubleComplex *) work, slab));
    return SQ_OK;
}

// out4[i] = {potential, fx, fy, fz} for the routed particles, from the four slabs with their ghost planes filled
int sq_pm_slab_readout(sq_ctx * c, const double * part4, int64_t m, int rank, int nranks, const double * s0, const double * s1,
                       const double * s2, const double * s3, double * out4)
{
    SQ_TRY(slab_check(c, nranks, "sq_pm_slab_readout"));
    const int N = c->par.Nmesh, nxl = N / nranks;
    if(m > 0) {
        sqd_slab_readout_kernel<<<dist_grid(c, m, 8), 256, 0, c->stream>>>((const double4 *) part4, m, c->par.BoxSize / N, N, rank * nxl, s0, s1, s2, s3, (double4 *) out4);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    return SQ_OK;
}

} /* extern "C" */
