/* TEST INFRASTRUCTURE ONLY - single-rank stand-in for <heffte.h>.
 *
 * The reference's PM (libgadget/petapm.cpp) runs its FFTs through heFFTe's fft3d_r2c with the FFTW backend hard-wired
 * (petapm.cpp:38, 298-300).  heFFTe needs cmake-generated configuration and FFTW3, neither of which this image has, so
 * the reference's PM could not be compiled in round 1.  This header provides just the part of the heFFTe API that
 * petapm.cpp uses (box3d, split_world, plan_options / default_options, fft3d_r2c::forward / backward / size_inbox /
 * size_outbox, scale) for ONE rank, with a plain double-precision mixed-radix FFT of our own behind it, so that the
 * reference's own paint, region, pencil-layout, transfer and read-out code (petapm.cpp, gravpm.cpp, powerspectrum.cpp)
 * compiles UNMODIFIED into oracle/_ref/ and pins the PM restatement of oracle/shenqi_oracle.c and tests/helpers.py.
 * Only the third-party FFT arithmetic is replaced; conventions are FFTW's / heFFTe's: forward = exp(-i...), backward =
 * exp(+i...), both unscaled with scale::none, c2r uses the real parts of the self-conjugate modes only.
 *
 * Written from the published API, not from the heFFTe sources.  Nothing in the product (shenqi_b200/) includes this.
 */
#ifndef SHENQI_ORACLE_HEFFTE_SHIM_H
#define SHENQI_ORACLE_HEFFTE_SHIM_H

#include <mpi.h>
#include <array>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace heffte {

template <typename index = int>
struct box3d {
    box3d(std::array<index, 3> l, std::array<index, 3> h) : box3d(l, h, {0, 1, 2}) {}
    box3d(std::array<index, 3> l, std::array<index, 3> h, std::array<int, 3> o)
        : low(l), high(h), size({h[0] - l[0] + 1, h[1] - l[1] + 1, h[2] - l[2] + 1}), order(o) {}
    bool empty() const { return size[0] <= 0 || size[1] <= 0 || size[2] <= 0; }
    long long count() const { return empty() ? 0 : (long long) size[0] * size[1] * size[2]; }
    /* the box shortened along `dimension` to the non-redundant half of a real-to-complex transform */
    box3d r2c(int dimension) const
    {
        std::array<index, 3> h = high;
        h[dimension] = low[dimension] + size[dimension] / 2;
        return box3d(low, h, order);
    }
    std::array<index, 3> low, high, size;
    std::array<int, 3> order;      /* order[0] = fastest dimension in memory */
};

/* block decomposition of `world` over a grid of tasks, first grid dimension varying fastest in the returned list */
inline std::vector<box3d<>> split_world(const box3d<> & world, std::array<int, 3> grid)
{
    std::vector<box3d<>> out;
    auto lo = [&](int d, int c) { return world.low[d] + (int) (((long long) world.size[d] * c) / grid[d]); };
    for(int k = 0; k < grid[2]; k++)
        for(int j = 0; j < grid[1]; j++)
            for(int i = 0; i < grid[0]; i++)
                out.push_back(box3d<>({lo(0, i), lo(1, j), lo(2, k)}, {lo(0, i + 1) - 1, lo(1, j + 1) - 1, lo(2, k + 1) - 1}, world.order));
    return out;
}

namespace backend { struct fftw {}; struct cufft {}; struct stock {}; }
enum class scale { none, full, symmetric };
struct plan_options { bool use_reorder = true; bool use_pencils = true; bool use_gpu_aware = false; };
template <class backend_tag> inline plan_options default_options() { return plan_options(); }

namespace shim {
typedef std::complex<double> cd;

/* 1-D complex DFT of any length: recursive decimation in time over the prime factors of n, twiddles from one table */
struct plan1d {
    int n = 0;
    std::vector<cd> w;      /* w[k] = exp(-2 pi i k / n) */
    explicit plan1d(int n_) : n(n_), w((size_t) n_)
    {
        for(int k = 0; k < n; k++) {
            const long double a = -2.0L * 3.141592653589793238462643383279502884L * (long double) k / (long double) n;
            w[k] = cd((double) cosl(a), (double) sinl(a));
        }
    }
    static int smallest_factor(int m)
    {
        for(int p = 2; (long long) p * p <= m; p++) if(m % p == 0) return p;
        return m;
    }
    cd tw(long long k, int sign) const { const cd v = w[(size_t) (k % n)]; return sign < 0 ? v : std::conj(v); }
    /* out[0 .. m) = DFT_m of in[0], in[istride], ...; m divides n; scratch holds at least (largest prime factor) entries */
    void rec(int m, const cd * in, long long istride, cd * out, int sign, cd * scratch) const
    {
        if(m == 1) { out[0] = in[0]; return; }
        const int p = smallest_factor(m), q = m / p;
        for(int r = 0; r < p; r++) rec(q, in + r * istride, istride * p, out + (size_t) r * q, sign, scratch);
        const long long step = n / m;     /* exp(-2 pi i / m) = w[step] */
        for(int k = 0; k < q; k++) {
            for(int r = 0; r < p; r++) scratch[r] = out[(size_t) r * q + k] * tw((long long) r * k * step, sign);
            for(int s = 0; s < p; s++) {
                cd acc = scratch[0];
                for(int r = 1; r < p; r++) acc += scratch[r] * tw((long long) r * s * q * step, sign);
                out[(size_t) s * q + k] = acc;
            }
        }
    }
    /* transform `line` (contiguous, n entries) in place; tmp: n + n entries of work space */
    void run(cd * line, int sign, cd * tmp) const
    {
        rec(n, line, 1, tmp, sign, tmp + n);
        for(int k = 0; k < n; k++) line[k] = tmp[k];
    }
};
}   /* namespace shim */

template <class backend_tag>
class fft3d_r2c {
public:
    fft3d_r2c(const box3d<> & inbox, const box3d<> & outbox, int r2c_direction, MPI_Comm comm, plan_options opts = plan_options())
        : in_(inbox), out_(outbox), pz_(inbox.size[2]), py_(inbox.size[1]), px_(inbox.size[0])
    {
        (void) opts;
        int ntask = 1;
        MPI_Comm_size(comm, &ntask);
        const bool whole = inbox.low[0] == 0 && inbox.low[1] == 0 && inbox.low[2] == 0 && outbox.low[0] == 0 && outbox.low[1] == 0 &&
                           outbox.low[2] == 0 && outbox.size[0] == inbox.size[0] && outbox.size[1] == inbox.size[1] &&
                           outbox.size[2] == inbox.size[2] / 2 + 1;
        if(ntask != 1 || r2c_direction != 2 || !whole) {
            fprintf(stderr, "oracle heffte shim: only one rank holding the whole mesh, r2c along z, is supported\n");
            abort();
        }
        strides(in_, si_);
        strides(out_, so_);
    }
    long long size_inbox() const { return in_.count(); }
    long long size_outbox() const { return out_.count(); }

    void forward(const double * real, std::complex<double> * cplx, scale s) const
    {
        const int nx = in_.size[0], ny = in_.size[1], nz = in_.size[2], nzc = nz / 2 + 1;
        std::vector<shim::cd> H((size_t) nx * ny * nzc);
        #pragma omp parallel
        {
            std::vector<shim::cd> line((size_t) nz), tmp(2 * (size_t) nz);
            #pragma omp for collapse(2)
            for(int x = 0; x < nx; x++)
                for(int y = 0; y < ny; y++) {
                    for(int z = 0; z < nz; z++) line[z] = real[x * si_[0] + y * si_[1] + z * si_[2]];
                    pz_.run(line.data(), -1, tmp.data());
                    for(int z = 0; z < nzc; z++) H[((size_t) x * ny + y) * nzc + z] = line[z];
                }
        }
        xy_passes(H, -1);
        const double f = factor(s);
        #pragma omp parallel for collapse(2)
        for(int x = 0; x < nx; x++)
            for(int y = 0; y < ny; y++)
                for(int z = 0; z < nzc; z++)
                    cplx[x * so_[0] + y * so_[1] + z * so_[2]] = f * H[((size_t) x * ny + y) * nzc + z];
    }

    void backward(const std::complex<double> * cplx, double * real, scale s) const
    {
        const int nx = in_.size[0], ny = in_.size[1], nz = in_.size[2], nzc = nz / 2 + 1;
        std::vector<shim::cd> H((size_t) nx * ny * nzc);
        #pragma omp parallel for collapse(2)
        for(int x = 0; x < nx; x++)
            for(int y = 0; y < ny; y++)
                for(int z = 0; z < nzc; z++)
                    H[((size_t) x * ny + y) * nzc + z] = cplx[x * so_[0] + y * so_[1] + z * so_[2]];
        xy_passes(H, +1);
        const double f = factor(s);
        #pragma omp parallel
        {
            std::vector<shim::cd> line((size_t) nz), tmp(2 * (size_t) nz);
            #pragma omp for collapse(2)
            for(int x = 0; x < nx; x++)
                for(int y = 0; y < ny; y++) {
                    const shim::cd * h = &H[((size_t) x * ny + y) * nzc];
                    /* Hermitian extension along z; the self-conjugate entries contribute their real parts only */
                    for(int z = 0; z < nzc; z++) line[z] = h[z];
                    for(int z = nzc; z < nz; z++) line[z] = std::conj(h[nz - z]);
                    line[0] = shim::cd(line[0].real(), 0.);
                    if(nz % 2 == 0) line[nz / 2] = shim::cd(line[nz / 2].real(), 0.);
                    pz_.run(line.data(), +1, tmp.data());
                    for(int z = 0; z < nz; z++) real[x * si_[0] + y * si_[1] + z * si_[2]] = f * line[z].real();
                }
        }
    }

private:
    static void strides(const box3d<> & b, long long * st)
    {
        st[b.order[0]] = 1;
        st[b.order[1]] = b.size[b.order[0]];
        st[b.order[2]] = (long long) b.size[b.order[0]] * b.size[b.order[1]];
    }
    double factor(scale s) const
    {
        const double n = (double) in_.count();
        return s == scale::none ? 1. : (s == scale::full ? 1. / n : 1. / std::sqrt(n));
    }
    /* complex transforms along y, then x, of the half mesh H[x][y][kz] */
    void xy_passes(std::vector<shim::cd> & H, int sign) const
    {
        const int nx = in_.size[0], ny = in_.size[1], nzc = in_.size[2] / 2 + 1;
        #pragma omp parallel
        {
            std::vector<shim::cd> line((size_t) (nx > ny ? nx : ny)), tmp(2 * (size_t) (nx > ny ? nx : ny));
            #pragma omp for collapse(2)
            for(int x = 0; x < nx; x++)
                for(int z = 0; z < nzc; z++) {
                    for(int y = 0; y < ny; y++) line[y] = H[((size_t) x * ny + y) * nzc + z];
                    py_.run(line.data(), sign, tmp.data());
                    for(int y = 0; y < ny; y++) H[((size_t) x * ny + y) * nzc + z] = line[y];
                }
            #pragma omp for collapse(2)
            for(int y = 0; y < ny; y++)
                for(int z = 0; z < nzc; z++) {
                    for(int x = 0; x < nx; x++) line[x] = H[((size_t) x * ny + y) * nzc + z];
                    px_.run(line.data(), sign, tmp.data());
                    for(int x = 0; x < nx; x++) H[((size_t) x * ny + y) * nzc + z] = line[x];
                }
        }
    }
    box3d<> in_, out_;
    shim::plan1d pz_, py_, px_;
    long long si_[3], so_[3];
};

}   /* namespace heffte */
#endif
