// Particle-mesh long-range force on one GPU: CIC paint -> r2c -> Green's function + differencing -> 4 x c2r
// -> fused CIC readout.
//
// Replaces, for the gravity callback set (libgadget/gravpm.cpp:32-39, 74):
//   pm_iterate_one + put_particle_to_mesh      libgadget/petapm.cpp:1132-1183, 1304-1310   (paint)
//   potential_transfer / force_*_transfer      libgadget/gravpm.cpp:378-487 via pm_apply_transfer_function petapm.cpp:1258-1298
//   petapm_fft_r2c / petapm_fft_c2r            libgadget/petapm.cpp:49-71 (heFFTe, unscaled both ways) -> cuFFT D2Z / Z2D
//   readout_potential / readout_force_{x,y,z}  libgadget/gravpm.cpp:489-500                (readout, 4 passes fused into 1)
// The reference's region/pencil exchange (petapm.cpp:692-1107) moves cells between the domain owner and
// the FFT slab owner; on one GPU the whole periodic mesh is one slab and cell indices wrap instead.
//
// Paint: particles arrive in Peano-Hilbert order, so a warp's 32 particles touch neighbouring cells;
// lanes that hit the same cell are combined with match/shuffle and ONE red.add.f64 is issued per distinct
// cell per warp (warp-aggregated atomics), the rest go to L2 as independent reductions.
// HBM-bound stages; algorithmic bytes in DESIGN.md.
#include "sq_common.cuh"
#include <math.h>

__device__ __forceinline__ int sq_wrap(int i, int n) { return i >= n ? i - n : (i < 0 ? i + n : i); }

// petapm.cpp:1146-1158: t = Pos/CellSize, iCell = floor(t), Res = t - iCell
__device__ __forceinline__ void sq_cic_cell(double p, double cell, int nmesh, int & i0, int & i1, double & res)
{
    const double t = p / cell;
    const double f = floor(t);
    res = t - f;
    int i = (int) f;
    i %= nmesh;
    if(i < 0) i += nmesh;
    i0 = i;
    i1 = sq_wrap(i + 1, nmesh);
}

__global__ void __launch_bounds__(256) sq_pm_paint_kernel(const double * __restrict__ pos, const float * __restrict__ mass, long long n,
                                                          double cell, int nmesh, double * __restrict__ mesh)
{
    const unsigned lane = sq_lane();
    const long long nwarp_iter = (n + 31) / 32;
    const long long warp0 = ((long long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long) gridDim.x * blockDim.x) >> 5;
    for(long long wi = warp0; wi < nwarp_iter; wi += nwarps) {
        const long long i = wi * 32 + lane;
        const bool ok = i < n;
        int ix[2] = {0, 0}, iy[2] = {0, 0}, iz[2] = {0, 0};
        double rx = 0, ry = 0, rz = 0, m = 0;
        if(ok) {
            sq_cic_cell(pos[3 * i], cell, nmesh, ix[0], ix[1], rx);
            sq_cic_cell(pos[3 * i + 1], cell, nmesh, iy[0], iy[1], ry);
            sq_cic_cell(pos[3 * i + 2], cell, nmesh, iz[0], iz[1], rz);
            m = (double) mass[i];
        }
        #pragma unroll
        for(int c = 0; c < 8; c++) {
            const int ox = c & 1, oy = (c >> 1) & 1, oz = (c >> 2) & 1;
            // petapm.cpp:1165-1176: weight = 1.0 * wx * wy * wz
            double w = 1.0;
            w *= ox ? rx : (1 - rx);
            w *= oy ? ry : (1 - ry);
            w *= oz ? rz : (1 - rz);
            const double v = w * m;
            const long long lin = ok ? ((long long) ix[ox] * nmesh + iy[oy]) * nmesh + iz[oz] : -1 - (long long) lane;
            // combine lanes that hit the same cell (hash = low bits of the three cell indices)
            const unsigned peers = sq_match64(lin, (unsigned) (ix[ox] & 7) | ((unsigned) (iy[oy] & 7) << 3) | ((unsigned) (iz[oz] & 3) << 6));
            unsigned rem = peers;
            double sum = 0;
            while(__any_sync(SQ_FULL, rem != 0)) {
                const int src = rem ? (__ffs(rem) - 1) : (int) lane;
                const double got = __shfl_sync(SQ_FULL, v, src);
                if(rem) { sum += got; rem &= rem - 1; }
            }
            if(ok && (peers & sq_lanemask_lt()) == 0)
                atomicAdd(&mesh[lin], sum);
        }
    }
}

// 4 fields fused: out[i] = {potential, fx, fy, fz}
__global__ void __launch_bounds__(256) sq_pm_readout_kernel(const double * __restrict__ pos, long long n, double cell, int nmesh,
                                                            const double * __restrict__ m0, const double * __restrict__ m1,
                                                            const double * __restrict__ m2, const double * __restrict__ m3,
                                                            double * __restrict__ out)
{
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        int ix[2], iy[2], iz[2];
        double rx, ry, rz;
        sq_cic_cell(pos[3 * i], cell, nmesh, ix[0], ix[1], rx);
        sq_cic_cell(pos[3 * i + 1], cell, nmesh, iy[0], iy[1], ry);
        sq_cic_cell(pos[3 * i + 2], cell, nmesh, iz[0], iz[1], rz);
        double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        #pragma unroll
        for(int c = 0; c < 8; c++) {
            const int ox = c & 1, oy = (c >> 1) & 1, oz = (c >> 2) & 1;
            double w = 1.0;
            w *= ox ? rx : (1 - rx);
            w *= oy ? ry : (1 - ry);
            w *= oz ? rz : (1 - rz);
            const long long lin = ((long long) ix[ox] * nmesh + iy[oy]) * nmesh + iz[oz];
            a0 += w * m0[lin];
            a1 += w * m1[lin];
            a2 += w * m2[lin];
            a3 += w * m3[lin];
        }
        out[4 * i] = a0; out[4 * i + 1] = a1; out[4 * i + 2] = a2; out[4 * i + 3] = a3;
    }
}

// One pass over the modes: phi_k from rho_k (gravpm.cpp:378-434), then the three force components from
// phi_k (gravpm.cpp:448-487).  isinc2[k] = 1/sinc^2(pi k/N), dker[k] = -diff_kernel(2 pi k/N) * N/Box.
__global__ void __launch_bounds__(256) sq_pm_transfer_kernel(const double2 * rho, int nmesh, double asmth2, double pot_factor,
                                                             const double * __restrict__ isinc2, const double * __restrict__ dker,
                                                             double2 * potk, double2 * __restrict__ fxk,
                                                             double2 * __restrict__ fyk, double2 * __restrict__ fzk)
{
    const int nz = nmesh / 2 + 1;
    const long long total = (long long) nmesh * nmesh * nz;
    for(long long ip = (long long) blockIdx.x * blockDim.x + threadIdx.x; ip < total; ip += (long long) gridDim.x * blockDim.x) {
        const int iz = (int) (ip % nz);
        const long long t = ip / nz;
        const int iy = (int) (t % nmesh);
        const int ix = (int) (t / nmesh);
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
        if(fxk) {       // k-space differencing (the reference's four c2r transforms); the product path differences phi in real space
            const double fx = dker[ix], fy = dker[iy], fz = dker[iz];
            fxk[ip] = make_double2(-v.y * fx, v.x * fx);
            fyk[ip] = make_double2(-v.y * fy, v.x * fy);
            fzk[ip] = make_double2(-v.y * fz, v.x * fz);
        }
    }
}

// Force meshes from the potential mesh by the 4-point central difference
//     F = -[ 2/3 (phi(+1) - phi(-1)) - 1/12 (phi(+2) - phi(-2)) ] N / Box
// which is the real-space form of the reference's k-space filter i K(w), K(w) = (8 sin w - sin 2w) / 6
// (gravpm.cpp:447-475: "the same as GADGET-2 but in fourier space").  One c2r transform instead of four; equal to the
// k-space result to rounding (tests: 1e-9 of the restatement, which keeps the four transforms).
// phi addresses plane 0 of nout output planes; xwrap: plane indices wrap mod nmesh (single GPU), otherwise planes
// -2 .. nout + 1 exist around it (x-slab with ghost planes).  y and z are periodic in the mesh.
//
// One thread owns a (y, z) column for GRAD_XCHUNK consecutive planes and marches along x with the five x-stencil values
// in registers, so the potential comes from DRAM once (the y and z neighbours of a plane are in flight in the
// neighbouring threads of the same plane: L1 / L2 hits).  The first version (one thread per cell, grid-stride over the
// volume) read the potential 4x from DRAM (ncu: 7.5 GB per pass for 4.3 GB of algorithmic traffic at 512^3).
#define GRAD_XCHUNK 32
__global__ void __launch_bounds__(256) sq_pm_gradient_kernel(const double * __restrict__ phi, int nmesh, int nout, int xwrap, double scale,
                                                             double * __restrict__ fx, double * __restrict__ fy, double * __restrict__ fz)
{
    const long long plane = (long long) nmesh * nmesh;
    const int nchunk = (nout + GRAD_XCHUNK - 1) / GRAD_XCHUNK;
    const long long total = plane * nchunk;
    const double c1 = -scale * (2.0 / 3.0), c2 = scale / 12.0;
    for(long long it = (long long) blockIdx.x * blockDim.x + threadIdx.x; it < total; it += (long long) gridDim.x * blockDim.x) {
        const long long yz = it % plane;            // consecutive threads: consecutive z
        const int chunk = (int) (it / plane);
        const int iz = (int) (yz % nmesh), iy = (int) (yz / nmesh);
        const int x0 = chunk * GRAD_XCHUNK;
        const int x1 = x0 + GRAD_XCHUNK < nout ? x0 + GRAD_XCHUNK : nout;
        const int ym2 = iy - 2 + (iy < 2 ? nmesh : 0), ym1 = iy - 1 + (iy < 1 ? nmesh : 0);
        const int yp1 = iy + 1 - (iy + 1 >= nmesh ? nmesh : 0), yp2 = iy + 2 - (iy + 2 >= nmesh ? nmesh : 0);
        const int zm2 = iz - 2 + (iz < 2 ? nmesh : 0), zm1 = iz - 1 + (iz < 1 ? nmesh : 0);
        const int zp1 = iz + 1 - (iz + 1 >= nmesh ? nmesh : 0), zp2 = iz + 2 - (iz + 2 >= nmesh ? nmesh : 0);
        const long long row = (long long) iy * nmesh;
        // plane index of x + d: periodic on one GPU; on an x-slab the planes -2 .. nout + 1 exist around plane 0
        auto px = [&](int x) -> long long {
            if(xwrap) { x += x < 0 ? nmesh : 0; x -= x >= nmesh ? nmesh : 0; }
            return (long long) x * plane;
        };
        double pm2 = phi[px(x0 - 2) + yz], pm1 = phi[px(x0 - 1) + yz], p0 = phi[px(x0) + yz], pp1 = phi[px(x0 + 1) + yz];
        for(int ix = x0; ix < x1; ix++) {
            const double pp2 = phi[px(ix + 2) + yz];
            const long long col = (long long) ix * plane;
            const long long ip = col + yz;
            fx[ip] = c1 * (pp1 - pm1) + c2 * (pp2 - pm2);
            fy[ip] = c1 * (phi[col + (long long) yp1 * nmesh + iz] - phi[col + (long long) ym1 * nmesh + iz])
                   + c2 * (phi[col + (long long) yp2 * nmesh + iz] - phi[col + (long long) ym2 * nmesh + iz]);
            fz[ip] = c1 * (phi[col + row + zp1] - phi[col + row + zm1]) + c2 * (phi[col + row + zp2] - phi[col + row + zm2]);
            pm2 = pm1; pm1 = p0; p0 = pp1; pp1 = pp2;
        }
    }
}

int sq_pm_gradient(sq_ctx * c, const double * phi, int nout, int xwrap, double * fx, double * fy, double * fz)
{
    const int N = c->par.Nmesh;
    const long long total = (long long) N * N * ((nout + GRAD_XCHUNK - 1) / GRAD_XCHUNK);
    long long need = (total + 255) / 256, cap = (long long) c->num_sms * 16;
    if(need < 1) need = 1;
    sq_pm_gradient_kernel<<<(int) (need < cap ? need : cap), 256, 0, c->stream>>>(phi, N, nout, xwrap, (double) N / c->par.BoxSize, fx, fy, fz);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

static double h_sinc_unnormed(double x)
{
    if(x < 1e-5 && x > -1e-5) {
        const double x2 = x * x;
        return 1.0 - x2 / 6. + x2 * x2 / 120.;
    }
    return sin(x) / x;
}

void sq_pm_destroy(sq_ctx * c)
{
    if(c->plan_r2c) cufftDestroy(c->plan_r2c);
    if(c->plan_c2r) cufftDestroy(c->plan_c2r);
    c->plan_r2c = c->plan_c2r = 0;
    c->plan_nmesh = 0;
}

static int pm_grid(const sq_ctx * c, long long n, int per_sm)
{
    long long need = (n + 255) / 256;
    long long cap = (long long) c->num_sms * per_sm;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}

int sq_pm_run(sq_ctx * c, bool paint_only)
{
    if(!c->sorted) { sq_set_error("sq_pm_force: call sq_ph_sort first"); return SQ_ERR_ARG; }
    const int N = c->par.Nmesh;
    if(N < 2 || (N & 1)) { sq_set_error("sq_pm_force: Nmesh must be even and >= 2"); return SQ_ERR_ARG; }
    const long long n = c->n_valid, nall = c->n;       // garbage particles neither paint nor read out
    const size_t nreal = (size_t) N * N * N;
    const size_t ncplx = (size_t) N * N * (N / 2 + 1);
    const double cell = c->par.BoxSize / N;
    SQ_TRY(c->pm_real[0].reserve(sizeof(double) * nreal));
    SQ_CUDA(cudaEventRecord(c->ev[10], c->stream));
    SQ_CUDA(cudaMemsetAsync(c->pm_real[0].p, 0, sizeof(double) * nreal, c->stream));
    if(n > 0) {
        sq_pm_paint_kernel<<<pm_grid(c, n, 8), 256, 0, c->stream>>>(c->pos.as<double>(), c->mass.as<float>(), n, cell, N, c->pm_real[0].as<double>());
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    SQ_CUDA(cudaEventRecord(c->ev[11], c->stream));
    if(paint_only) {
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0; cudaEventElapsedTime(&ms, c->ev[10], c->ev[11]); c->tm.pm_paint_ms = ms;
        return SQ_OK;
    }
    SQ_TRY(c->pm_cplx[0].reserve(sizeof(double2) * ncplx));
    for(int f = 1; f < 4; f++) SQ_TRY(c->pm_real[f].reserve(sizeof(double) * nreal));
    SQ_TRY(c->pm_out.reserve(sizeof(double) * 4 * (size_t) (nall + 1) + sizeof(double) * 2 * (size_t) N));
    if(nall > n) SQ_CUDA(cudaMemsetAsync(c->pm_out.as<double>() + 4 * (size_t) n, 0, sizeof(double) * 4 * (size_t) (nall - n), c->stream));
    if(c->plan_nmesh != N) {
        sq_pm_destroy(c);
        SQ_CUFFT(cufftPlan3d(&c->plan_r2c, N, N, N, CUFFT_D2Z));
        SQ_CUFFT(cufftPlan3d(&c->plan_c2r, N, N, N, CUFFT_Z2D));
        SQ_CUFFT(cufftSetStream(c->plan_r2c, c->stream));
        SQ_CUFFT(cufftSetStream(c->plan_c2r, c->stream));
        c->plan_nmesh = N;
    }
    // per-axis factors: gravpm.cpp:395-401 (CIC deconvolution) and :436-460 (differencing)
    std::vector<double> tab(2 * (size_t) N);
    for(int i = 0; i < N; i++) {
        const int k = i <= N / 2 ? i : i - N;
        double tmp = (k * M_PI) / N;
        tmp = h_sinc_unnormed(tmp);
        tab[i] = 1. / (tmp * tmp);
        const double w = k * (2 * M_PI / N);
        tab[N + i] = -1 * (1 / 6.0 * (8 * sin(w) - sin(2 * w))) * (N / c->par.BoxSize);
    }
    double * d_tab = c->pm_out.as<double>() + 4 * (size_t) (nall + 1);
    SQ_CUDA(cudaMemcpyAsync(d_tab, tab.data(), sizeof(double) * 2 * (size_t) N, cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));     // tab is a host temporary
    SQ_CUFFT(cufftExecD2Z(c->plan_r2c, c->pm_real[0].as<double>(), c->pm_cplx[0].as<cufftDoubleComplex>()));
    SQ_CUDA(cudaEventRecord(c->ev[12], c->stream));
    const double asmth2 = pow((2 * M_PI) * c->par.Asmth / N, 2);
    const double pot_factor = -c->par.G / (M_PI * c->par.BoxSize);
    sq_pm_transfer_kernel<<<pm_grid(c, (long long) ncplx, 16), 256, 0, c->stream>>>(c->pm_cplx[0].as<double2>(), N, asmth2, pot_factor, d_tab, d_tab + N,
            c->pm_cplx[0].as<double2>(), nullptr, nullptr, nullptr);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    SQ_CUDA(cudaEventRecord(c->ev[13], c->stream));
    SQ_CUFFT(cufftExecZ2D(c->plan_c2r, c->pm_cplx[0].as<cufftDoubleComplex>(), c->pm_real[0].as<double>()));
    SQ_TRY(sq_pm_gradient(c, c->pm_real[0].as<double>(), N, 1, c->pm_real[1].as<double>(), c->pm_real[2].as<double>(), c->pm_real[3].as<double>()));
    SQ_CUDA(cudaEventRecord(c->ev[14], c->stream));
    if(n > 0) {
        sq_pm_readout_kernel<<<pm_grid(c, n, 8), 256, 0, c->stream>>>(c->pos.as<double>(), n, cell, N, c->pm_real[0].as<double>(), c->pm_real[1].as<double>(),
                c->pm_real[2].as<double>(), c->pm_real[3].as<double>(), c->pm_out.as<double>());
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    SQ_CUDA(cudaEventRecord(c->ev[15], c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[10], c->ev[11]); c->tm.pm_paint_ms = ms;
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, c->ev[11], c->ev[12]);
    cudaEventElapsedTime(&b, c->ev[13], c->ev[14]);
    c->tm.pm_fft_ms = a + b;
    cudaEventElapsedTime(&ms, c->ev[12], c->ev[13]); c->tm.pm_transfer_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[14], c->ev[15]); c->tm.pm_readout_ms = ms;
    c->pm_done = true;
    return SQ_OK;
}
