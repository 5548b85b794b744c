// Stable LSD radix sort of (64-bit Peano-Hilbert key, 32-bit index) pairs, and the exclusive scan
// it is built on.  Replaces std::sort over struct PeanoOrder (libgadget/slotsmanager.cpp:395-437);
// ties keep upload order (the reference's unstable sort leaves them undefined).
//
// 8-bit digits.  Per pass: (1) per-tile digit histogram, (2) exclusive scan of the bin-major
// histogram matrix, (3) re-read the tile, rank every key inside the block with warp match/ballot
// (eight ballots per key - MATCH.ANY is slow on sm_100 -, stable), stage the tile in shared memory in digit order and write each digit run
// out contiguously.  Algorithmic traffic per pass: 8 B (histogram read) + 12 B read + 12 B written.
#include "sq_common.cuh"

#define SORT_THREADS 256
#define SORT_ITEMS 12
#define SORT_TILE (SORT_THREADS * SORT_ITEMS)     /* 3072 pairs: 36 KB staged */
#define SORT_WARPS (SORT_THREADS / 32)

// ------------------------------------------------------------------------------------ scan
#define SCAN_THREADS 256
#define SCAN_ITEMS 8
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

// in-place exclusive scan of one tile per block; block totals to sums[block] (if sums != NULL)
__global__ void __launch_bounds__(SCAN_THREADS) sq_scan_tile_kernel(uint32_t * __restrict__ data, int64_t n, uint32_t * __restrict__ sums)
{
    __shared__ uint32_t warp_tot[SCAN_THREADS / 32];
    const int64_t base = (int64_t) blockIdx.x * SCAN_TILE + (int64_t) threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t tsum = 0;
    #pragma unroll
    for(int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? data[base + k] : 0u;
        tsum += v[k];
    }
    int total;
    int excl = sq_warp_excl_scan((int) tsum, total);
    const int w = threadIdx.x >> 5;
    if(sq_lane() == 31) warp_tot[w] = (uint32_t) total;
    __syncthreads();
    uint32_t woff = 0, btot = 0;
    #pragma unroll
    for(int k = 0; k < SCAN_THREADS / 32; k++) {
        if(k < w) woff += warp_tot[k];
        btot += warp_tot[k];
    }
    uint32_t run = woff + (uint32_t) excl;
    #pragma unroll
    for(int k = 0; k < SCAN_ITEMS; k++) {
        if(base + k < n) data[base + k] = run;
        run += v[k];
    }
    if(sums && threadIdx.x == 0) sums[blockIdx.x] = btot;
}

__global__ void __launch_bounds__(SCAN_THREADS) sq_scan_add_kernel(uint32_t * __restrict__ data, int64_t n, const uint32_t * __restrict__ sums)
{
    const uint32_t add = sums[blockIdx.x];
    const int64_t base = (int64_t) blockIdx.x * SCAN_TILE;
    for(int k = threadIdx.x; k < SCAN_TILE; k += SCAN_THREADS)
        if(base + k < n) data[base + k] += add;
}

static int sq_scan_rec(sq_ctx * c, uint32_t * data, int64_t n, uint32_t * tmp)
{
    const int64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    if(nb <= 1) {
        sq_scan_tile_kernel<<<1, SCAN_THREADS, 0, c->stream>>>(data, n, nullptr);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
        return SQ_OK;
    }
    sq_scan_tile_kernel<<<(unsigned) nb, SCAN_THREADS, 0, c->stream>>>(data, n, tmp);
    c->kernel_launches++;
    SQ_TRY(sq_scan_rec(c, tmp, nb, tmp + ((nb + 63) & ~(int64_t) 63)));
    sq_scan_add_kernel<<<(unsigned) nb, SCAN_THREADS, 0, c->stream>>>(data, n, tmp);
    c->kernel_launches++;
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

__global__ void sq_last_plus_kernel(const uint32_t * scanned_last, const uint32_t * orig_last_saved, uint32_t * out)
{
    *out = *scanned_last + *orig_last_saved;
}

// Exclusive scan in place.  If total_out_dev != NULL, *total_out_dev = sum of all inputs.
int sq_exclusive_scan_u32(sq_ctx * c, uint32_t * data, int64_t n, uint32_t * total_out_dev)
{
    if(n <= 0) {
        if(total_out_dev) SQ_CUDA(cudaMemsetAsync(total_out_dev, 0, 4, c->stream));
        return SQ_OK;
    }
    size_t tmp_elems = 64;
    for(int64_t m = n; m > 1; ) { m = (m + SCAN_TILE - 1) / SCAN_TILE; tmp_elems += (size_t) ((m + 63) & ~(int64_t) 63); }
    SQ_TRY(c->scan_tmp.reserve(sizeof(uint32_t) * (tmp_elems + 64)));
    uint32_t * tmp = c->scan_tmp.as<uint32_t>();
    uint32_t * saved_last = tmp;      // first 64 entries reserved
    if(total_out_dev)
        SQ_CUDA(cudaMemcpyAsync(saved_last, data + (n - 1), 4, cudaMemcpyDeviceToDevice, c->stream));
    SQ_TRY(sq_scan_rec(c, data, n, tmp + 64));
    if(total_out_dev) {
        sq_last_plus_kernel<<<1, 1, 0, c->stream>>>(data + (n - 1), saved_last, total_out_dev);
        c->kernel_launches++;
    }
    return SQ_OK;
}

// ------------------------------------------------------------------------------------ radix passes
__device__ __forceinline__ unsigned sq_digit(uint64_t key, int shift) { return (unsigned) (key >> shift) & 255u; }

__global__ void __launch_bounds__(SORT_THREADS) sq_sort_hist_kernel(const uint64_t * __restrict__ keys, int64_t n, int shift,
                                                                    uint32_t * __restrict__ hist, int nblocks)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t) blockIdx.x * SORT_TILE;
    #pragma unroll
    for(int k = 0; k < SORT_ITEMS; k++) {
        const int64_t i = base + k * SORT_THREADS + threadIdx.x;
        if(i < n) {
            // warp-aggregate equal digits before touching shared memory
            const unsigned d = sq_digit(keys[i], shift);
            const unsigned active = __activemask();
            const unsigned peers = sq_match8(active, d);
            if((peers & sq_lanemask_lt()) == 0) atomicAdd(&h[d], (uint32_t) __popc(peers));
        }
    }
    __syncthreads();
    hist[(size_t) threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

struct __align__(16) sq_sort_smem {
    uint64_t skeys[SORT_TILE];
    uint32_t svals[SORT_TILE];
    uint32_t warp_hist[SORT_WARPS][256];
    uint32_t bin_start[256];
    uint32_t glob_off[256];
};

__global__ void __launch_bounds__(SORT_THREADS) sq_sort_scatter_kernel(const uint64_t * __restrict__ keys_in, const uint32_t * __restrict__ vals_in,
                                                                       uint64_t * __restrict__ keys_out, uint32_t * __restrict__ vals_out,
                                                                       int64_t n, int shift, const uint32_t * __restrict__ hist, int nblocks)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    sq_sort_smem & s = *reinterpret_cast<sq_sort_smem *>(smem_raw);
    const int w = threadIdx.x >> 5;
    const unsigned lane = sq_lane();
    for(int k = threadIdx.x; k < SORT_WARPS * 256; k += SORT_THREADS)
        (&s.warp_hist[0][0])[k] = 0;
    s.glob_off[threadIdx.x] = hist[(size_t) threadIdx.x * nblocks + blockIdx.x];
    __syncthreads();

    // warp w owns the contiguous elements [w*ITEMS*32, (w+1)*ITEMS*32) of the tile; round r, lane l
    const int64_t base = (int64_t) blockIdx.x * SORT_TILE + (int64_t) w * SORT_ITEMS * 32;
    uint64_t key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
    uint32_t rnk[SORT_ITEMS];
    #pragma unroll
    for(int r = 0; r < SORT_ITEMS; r++) {
        const int64_t i = base + r * 32 + lane;
        const bool ok = i < n;
        key[r] = ok ? keys_in[i] : ~(uint64_t) 0;
        val[r] = ok ? vals_in[i] : 0u;
    }
    #pragma unroll
    for(int r = 0; r < SORT_ITEMS; r++) {
        const int64_t i = base + r * 32 + lane;
        const bool ok = i < n;
        const unsigned valid = __ballot_sync(SQ_FULL, ok);
        rnk[r] = 0;
        if(ok) {
            const unsigned d = sq_digit(key[r], shift);
            const unsigned peers = sq_match8(valid, d);
            const int leader = __ffs(peers) - 1;
            uint32_t b = 0;
            // one shared-memory atomic per digit group
            if((int) lane == leader) b = atomicAdd(&s.warp_hist[w][d], (uint32_t) __popc(peers));
            b = __shfl_sync(peers, b, leader);
            rnk[r] = b + (uint32_t) __popc(peers & sq_lanemask_lt());
        }
        // stability needs round r's atomics applied before round r + 1's (different leader lanes may hit the same
        // digit): the barrier orders the warp's shared-memory accesses explicitly instead of relying on issue order
        __syncwarp();
    }
    __syncthreads();
    // thread d: exclusive scan of bin d over warps, then block-wide exclusive scan over bins
    {
        const int d = threadIdx.x;
        uint32_t run = 0;
        #pragma unroll
        for(int ww = 0; ww < SORT_WARPS; ww++) {
            const uint32_t t = s.warp_hist[ww][d];
            s.warp_hist[ww][d] = run;
            run += t;
        }
        int wtot;
        int excl = sq_warp_excl_scan((int) run, wtot);
        __shared__ uint32_t wsum[SORT_WARPS];
        if(lane == 31) wsum[w] = (uint32_t) wtot;
        __syncthreads();
        uint32_t off = 0;
        #pragma unroll
        for(int ww = 0; ww < SORT_WARPS; ww++)
            if(ww < w) off += wsum[ww];
        s.bin_start[d] = off + (uint32_t) excl;
    }
    __syncthreads();
    #pragma unroll
    for(int r = 0; r < SORT_ITEMS; r++) {
        const int64_t i = base + r * 32 + lane;
        if(i < n) {
            const unsigned d = sq_digit(key[r], shift);
            const uint32_t p = s.bin_start[d] + s.warp_hist[w][d] + rnk[r];
            s.skeys[p] = key[r];
            s.svals[p] = val[r];
        }
    }
    __syncthreads();
    const int64_t tile_base = (int64_t) blockIdx.x * SORT_TILE;
    const int cnt = (int) ((n - tile_base) < SORT_TILE ? (n - tile_base) : SORT_TILE);
    for(int p = threadIdx.x; p < cnt; p += SORT_THREADS) {
        const uint64_t k = s.skeys[p];
        const unsigned d = sq_digit(k, shift);
        const int64_t g = (int64_t) s.glob_off[d] + (p - (int) s.bin_start[d]);
        keys_out[g] = k;
        vals_out[g] = s.svals[p];
    }
}

// Sort pairs in place (result back in keys/vals) on bits [begin_bit, end_bit).
int sq_sort_pairs(sq_ctx * c, uint64_t * keys, uint32_t * vals, int64_t n, int begin_bit, int end_bit)
{
    if(n <= 1) return SQ_OK;
    if(n >= ((int64_t) 1 << 32)) { sq_set_error("sq_sort_pairs: n >= 2^32"); return SQ_ERR_ARG; }
    const int nblocks = (int) ((n + SORT_TILE - 1) / SORT_TILE);
    SQ_TRY(c->sort_tmp_keys.reserve(sizeof(uint64_t) * (size_t) n));
    SQ_TRY(c->sort_tmp_vals.reserve(sizeof(uint32_t) * (size_t) n));
    SQ_TRY(c->sort_hist.reserve(sizeof(uint32_t) * 256 * (size_t) nblocks));
    // per device, not per process: set it on every call (a cheap driver query) rather than behind a process-wide flag
    SQ_CUDA(cudaFuncSetAttribute(sq_sort_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sizeof(sq_sort_smem)));
    uint64_t * kin = keys, * kout = c->sort_tmp_keys.as<uint64_t>();
    uint32_t * vin = vals, * vout = c->sort_tmp_vals.as<uint32_t>();
    uint32_t * hist = c->sort_hist.as<uint32_t>();
    for(int shift = begin_bit; shift < end_bit; shift += 8) {
        sq_sort_hist_kernel<<<nblocks, SORT_THREADS, 0, c->stream>>>(kin, n, shift, hist, nblocks);
        c->kernel_launches++;
        SQ_TRY(sq_exclusive_scan_u32(c, hist, (int64_t) 256 * nblocks, nullptr));
        sq_sort_scatter_kernel<<<nblocks, SORT_THREADS, sizeof(sq_sort_smem), c->stream>>>(kin, vin, kout, vout, n, shift, hist, nblocks);
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
        uint64_t * tk = kin; kin = kout; kout = tk;
        uint32_t * tv = vin; vin = vout; vout = tv;
    }
    if(kin != keys) {
        SQ_CUDA(cudaMemcpyAsync(keys, kin, sizeof(uint64_t) * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
        SQ_CUDA(cudaMemcpyAsync(vals, vin, sizeof(uint32_t) * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
    }
    return SQ_OK;
}
