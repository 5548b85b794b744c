// Peano-Hilbert keys and the gather into key order.
//
// Replaces PEANO()/peano_hilbert_key() (libgadget/utils/peano.h:15-21, peano.cpp:144-183) and the
// permutation half of slots_gc_sorted (libgadget/slotsmanager.cpp:438-452).
//
// One thread per particle.  The 48x8 orientation tables are composed two levels at a time into
// 48x64 tables (6 KB) held in shared memory, so a 21-level key is 1 + 10 dependent shared-memory
// lookups.  HBM-bound: 24 B read + 8 B written per particle.
//
// Besides the key, every thread re-derives its octant path geometrically, with the reference's own
// FP64 recurrence (centre +- len/4, `Pos > centre`, libgadget/forcetree.cpp:277-283, 304-319), and
// raises a flag if the path differs from the key digits (it does for particles exactly on a cell
// face, e.g. lattices).  The tree build uses that flag to fall back to geometric keys, so the tree
// topology is the reference's even where key truncation and the geometric test disagree.
#include "sq_common.cuh"
#include "ph_tables.h"

static const unsigned char h_ph_rot[48 * 8] = SQ_PH_ROT_INIT;
static const unsigned char h_ph_sub[48 * 8] = SQ_PH_SUB_INIT;

// [0,3072): rot2[r*64+p]   [3072,6144): sub2[r*64+p]   [6144,6528): rot1[r*8+p]   [6528,6912): sub1
#define SQ_PHTAB_BYTES 6912
// one copy per device (a process may drive several GPUs through several contexts), created under a lock
#include <mutex>
static unsigned char * d_phtab_dev[64] = {};
static std::mutex d_phtab_lock;
#define d_phtab (d_phtab_dev[sq_current_device()])
static int sq_current_device() { int d = 0; cudaGetDevice(&d); return (d >= 0 && d < 64) ? d : 0; }

static int sq_phtab_init()
{
    std::lock_guard<std::mutex> guard(d_phtab_lock);
    if(d_phtab) return SQ_OK;
    unsigned char h[SQ_PHTAB_BYTES];
    for(int r = 0; r < 48; r++)
        for(int p = 0; p < 64; p++) {
            const int hi = p >> 3, lo = p & 7;
            const int r1 = h_ph_rot[r * 8 + hi];
            h[r * 64 + p] = h_ph_rot[r1 * 8 + lo];
            h[3072 + r * 64 + p] = (unsigned char) ((h_ph_sub[r * 8 + hi] << 3) | h_ph_sub[r1 * 8 + lo]);
        }
    memcpy(h + 6144, h_ph_rot, 384);
    memcpy(h + 6528, h_ph_sub, 384);
    SQ_CUDA(cudaMalloc((void **) &d_phtab, SQ_PHTAB_BYTES));
    SQ_CUDA(cudaMemcpy(d_phtab, h, SQ_PHTAB_BYTES, cudaMemcpyHostToDevice));
    return SQ_OK;
}

const unsigned char * sq_phtab_device()
{
    return sq_phtab_init() == SQ_OK ? d_phtab : nullptr;
}

__device__ __forceinline__ uint64_t sq_ph_from_ints(const unsigned char * tab, unsigned x, unsigned y, unsigned z)
{
    // level 20 alone (21 is odd), then two levels per lookup: peano.cpp:160-181
    unsigned pix = (((x >> 20) & 1) << 2) | (((y >> 20) & 1) << 1) | ((z >> 20) & 1);
    uint64_t key = tab[6528 + pix];
    unsigned rot = tab[6144 + pix];
    #pragma unroll
    for(int bit = 19; bit >= 1; bit -= 2) {
        const unsigned hi = (((x >> bit) & 1) << 2) | (((y >> bit) & 1) << 1) | ((z >> bit) & 1);
        const unsigned lo = (((x >> (bit - 1)) & 1) << 2) | (((y >> (bit - 1)) & 1) << 1) | ((z >> (bit - 1)) & 1);
        const unsigned p = rot * 64 + ((hi << 3) | lo);
        key = (key << 6) | tab[3072 + p];
        rot = tab[p];
    }
    return key;
}

// 21 geometric octant bits of one coordinate: forcetree.cpp:277-283 + 304-319 restated as a recurrence
__device__ __forceinline__ unsigned sq_geo_bits(double p, double box)
{
    double c = box / 2.;
    double q = 0.25 * (box * 1.001);
    unsigned g = 0;
    #pragma unroll
    for(int l = 0; l < SQ_MAXLEVEL; l++) {
        const bool up = p > c;
        g = (g << 1) | (up ? 1u : 0u);
        c = up ? __dadd_rn(c, q) : __dadd_rn(c, -q);
        q *= 0.5;
    }
    return g;
}

template <bool GEO>
__global__ void __launch_bounds__(256) sq_keys_kernel(const double * __restrict__ pos, int64_t n, double box, double box2000,
                                                      double domainfac, const unsigned char * __restrict__ phtab,
                                                      uint64_t * __restrict__ keys, int * __restrict__ flags)
{
    __shared__ __align__(16) unsigned char tab[SQ_PHTAB_BYTES];
    for(int i = threadIdx.x; i < SQ_PHTAB_BYTES / 16; i += blockDim.x)
        ((uint4 *) tab)[i] = ((const uint4 *) phtab)[i];
    __syncthreads();
    int bad = 0;
    for(int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double px = pos[3 * i], py = pos[3 * i + 1], pz = pos[3 * i + 2];
        const unsigned gx = sq_geo_bits(px, box), gy = sq_geo_bits(py, box), gz = sq_geo_bits(pz, box);
        if(GEO) {
            keys[i] = sq_ph_from_ints(tab, gx, gy, gz);
        } else {
            // peano.h:18-20: no contraction, truncation toward zero
            const unsigned ix = (unsigned) __double2int_rz(__dmul_rn(__dadd_rn(px, box2000), domainfac));
            const unsigned iy = (unsigned) __double2int_rz(__dmul_rn(__dadd_rn(py, box2000), domainfac));
            const unsigned iz = (unsigned) __double2int_rz(__dmul_rn(__dadd_rn(pz, box2000), domainfac));
            keys[i] = sq_ph_from_ints(tab, ix, iy, iz);
            bad |= (ix != gx) | (iy != gy) | (iz != gz);
        }
    }
    if(!GEO && __any_sync(SQ_FULL, bad) && sq_lane() == 0)
        atomicOr(flags, 1);
}

// garbage / swallowed particles: the largest key, so that they sort behind every valid particle
__global__ void sq_keys_skip_kernel(const unsigned char * __restrict__ skip, int64_t n, uint64_t * __restrict__ keys)
{
    for(int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x)
        if(skip[i]) keys[i] = ~0ull;
}

// perm[i] = upload index of the particle at sorted position i
__global__ void __launch_bounds__(256) sq_gather_kernel(const uint32_t * __restrict__ perm, int64_t n,
                                                        const double * __restrict__ pos_in, const float * __restrict__ mass_in,
                                                        const double * __restrict__ oldacc_in,
                                                        double * __restrict__ pos, float * __restrict__ mass,
                                                        double * __restrict__ oldacc, uint32_t * __restrict__ rank,
                                                        int * __restrict__ flags, int64_t nvalid)
{
    int bad = 0;
    for(int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const uint32_t j = perm[i];
        pos[3 * i] = pos_in[3 * (int64_t) j];
        pos[3 * i + 1] = pos_in[3 * (int64_t) j + 1];
        pos[3 * i + 2] = pos_in[3 * (int64_t) j + 2];
        const float m = mass_in[j];
        mass[i] = m;
        bad |= (i < nvalid) && !(m > 0.f);
        oldacc[i] = oldacc_in ? oldacc_in[j] : 0.0;
        rank[j] = (uint32_t) i;
    }
    if(__any_sync(SQ_FULL, bad) && sq_lane() == 0)
        atomicOr(flags, 2);   // Mass <= 0: forcetree.cpp:808
}

__global__ void sq_iota_kernel(uint32_t * v, int64_t n)
{
    for(int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x)
        v[i] = (uint32_t) i;
}

static int sq_grid_for(const sq_ctx * c, int64_t n, int threads, int per_sm)
{
    int64_t need = (n + threads - 1) / threads;
    int64_t cap = (int64_t) c->num_sms * per_sm;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}

int sq_geokeys_run(sq_ctx * c, const double * pos, int64_t n, uint64_t * gkeys)
{
    SQ_TRY(sq_phtab_init());
    if(n == 0) return SQ_OK;
    sq_keys_kernel<true><<<sq_grid_for(c, n, 256, 8), 256, 0, c->stream>>>(pos, n, c->par.BoxSize, 0, 0, d_phtab, gkeys, nullptr);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

// keys of the uploaded particles -> c->keys_in, geometric-mismatch flag -> c->geo_mismatch
int sq_keys_run(sq_ctx * c)
{
    SQ_TRY(sq_phtab_init());
    const int64_t n = c->n;
    SQ_TRY(c->keys_in.reserve(sizeof(uint64_t) * (size_t) (n + 1)));
    SQ_TRY(c->t_counters.reserve(256));
    SQ_CUDA(cudaMemsetAsync(c->t_counters.p, 0, 256, c->stream));
    if(n == 0) { c->geo_mismatch = 0; return SQ_OK; }
    const double box = c->par.BoxSize;
    // exactly the reference's expressions (peano.h:18-19)
    const double domainfac = 1.0 / (box * 1.001) * (double) (((uint64_t) 1) << 21);
    const double box2000 = box / 2000;
    SQ_CUDA(cudaEventRecord(c->ev[0], c->stream));
    sq_keys_kernel<false><<<sq_grid_for(c, n, 256, 8), 256, 0, c->stream>>>(c->pos_in.as<double>(), n, box, box2000, domainfac,
                                                                             d_phtab, c->keys_in.as<uint64_t>(), c->t_counters.as<int>());
    c->kernel_launches++;
    if(c->have_skip) {
        sq_keys_skip_kernel<<<sq_grid_for(c, n, 256, 8), 256, 0, c->stream>>>(c->skip_in.as<unsigned char>(), n, c->keys_in.as<uint64_t>());
        c->kernel_launches++;
    }
    SQ_CUDA(cudaGetLastError());
    SQ_CUDA(cudaEventRecord(c->ev[1], c->stream));
    return SQ_OK;
}

// sort (key,idx), gather particle fields into key order
int sq_gather_run(sq_ctx * c)
{
    const int64_t n = c->n;
    SQ_TRY(c->pos.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
    SQ_TRY(c->mass.reserve(sizeof(float) * (size_t) (n + 1)));
    SQ_TRY(c->oldacc.reserve(sizeof(double) * (size_t) (n + 1)));
    SQ_TRY(c->rank.reserve(sizeof(uint32_t) * (size_t) (n + 1)));
    if(n == 0) return SQ_OK;
    sq_gather_kernel<<<sq_grid_for(c, n, 256, 8), 256, 0, c->stream>>>(c->perm.as<uint32_t>(), n, c->pos_in.as<double>(),
            c->mass_in.as<float>(), c->oldacc_in.p ? c->oldacc_in.as<double>() : nullptr, c->pos.as<double>(), c->mass.as<float>(),
            c->oldacc.as<double>(), c->rank.as<uint32_t>(), c->t_counters.as<int>(), c->n_valid);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

int sq_iota(sq_ctx * c, uint32_t * v, int64_t n)
{
    if(n == 0) return SQ_OK;
    sq_iota_kernel<<<sq_grid_for(c, n, 256, 8), 256, 0, c->stream>>>(v, n);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

// PEANO() keys of positions that are already on the device (no context state touched except the parameter BoxSize)
extern "C" int sq_keys_device(sq_ctx * c, const double * pos, int64_t n, uint64_t * keys)
{
    if(!c || !c->have_params || (n > 0 && (!pos || !keys))) { sq_set_error("sq_keys_device: bad argument (set the parameters first)"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaSetDevice(c->device));
    SQ_TRY(sq_phtab_init());
    if(n == 0) return SQ_OK;
    SQ_TRY(c->t_counters.reserve(256));
    const double box = c->par.BoxSize;
    const double domainfac = 1.0 / (box * 1.001) * (double) (((uint64_t) 1) << 21);
    sq_keys_kernel<false><<<sq_grid_for(c, n, 256, 8), 256, 0, c->stream>>>(pos, n, box, box / 2000, domainfac, d_phtab, keys, c->t_counters.as<int>() + 8);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}
