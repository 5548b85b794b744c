// C ABI of libshenqi_b200.so (include/shenqi_b200.h).  Thin: argument checks, host<->device copies,
// stage orchestration.  No CPU fallback anywhere: every entry point needs the CUDA context.
#include "sq_common.cuh"
#include <stdarg.h>
#include <math.h>

static thread_local char sq_err[1024] = "";

void sq_set_error(const char * fmt, ...)
{
    va_list va;
    va_start(va, fmt);
    vsnprintf(sq_err, sizeof(sq_err), fmt, va);
    va_end(va);
}

int sq_iota(sq_ctx * c, uint32_t * v, int64_t n);

// out[perm[i]*K + k] = in[i*K + k]   (sorted order -> upload order)
template <int K>
__global__ void sq_unpermute_kernel(const double * __restrict__ in, int in_stride, int in_off, const uint32_t * __restrict__ perm,
                                    const uint32_t * __restrict__ targets, long long m, double * __restrict__ out)
{
    for(long long q = (long long) blockIdx.x * blockDim.x + threadIdx.x; q < m; q += (long long) gridDim.x * blockDim.x) {
        const long long i = targets ? targets[q] : q;
        const long long j = perm[i];
        #pragma unroll
        for(int k = 0; k < K; k++)
            out[j * K + k] = in[i * in_stride + in_off + k];
    }
}

// |FullTreeGravAccel + GravPM| / G, gravshort2.hpp:111-121
__global__ void sq_oldacc_kernel(const double * __restrict__ a, const double * __restrict__ b, long long n, double G, double * __restrict__ out)
{
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        double s = 0;
        for(int j = 0; j < 3; j++) {
            const double ax = (a ? a[3 * i + j] : 0.0) + (b ? b[3 * i + j] : 0.0);
            s += ax * ax;
        }
        out[i] = sqrt(s) / G;
    }
}

// unpack the caller's array-of-structs records (struct particle_data, partmanager.h:9-71)
__global__ void sq_unpack_aos_kernel(const unsigned char * __restrict__ base, long long n, sq_aos_layout lay, double G,
                                     double * __restrict__ pos, float * __restrict__ mass, double * __restrict__ oldacc,
                                     unsigned char * __restrict__ skip, int * __restrict__ nskip)
{
    for(long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long) gridDim.x * blockDim.x) {
        const unsigned char * p = base + (size_t) i * lay.stride;
        const double * P = (const double *) (p + lay.off_pos);
        pos[3 * i] = P[0]; pos[3 * i + 1] = P[1]; pos[3 * i + 2] = P[2];
        mass[i] = *(const float *) (p + lay.off_mass);
        const double * A = (const double *) (p + lay.off_fulltreegravaccel);
        const double * B = (const double *) (p + lay.off_gravpm);
        double s = 0;
        for(int j = 0; j < 3; j++) { const double ax = A[j] + B[j]; s += ax * ax; }
        oldacc[i] = sqrt(s) / G;
        // partmanager.h:18-31: bit 0 IsGarbage, bit 1 Swallowed, byte + 3 Type.  Garbage and swallowed black holes are not in
        // the tree (forcetree.cpp:805) and sort behind everything else (TypeKey 255, slotsmanager.cpp:428-432)
        const unsigned char fl = p[lay.off_flags], type = p[lay.off_flags + 3];
        const bool sk = (fl & 1u) || ((fl & 2u) && type == 5);
        skip[i] = sk ? 1 : 0;
        if(sk) atomicAdd(nskip, 1);
    }
}

static int api_grid(const sq_ctx * c, long long n)
{
    long long need = (n + 255) / 256, cap = (long long) c->num_sms * 8;
    if(need < 1) need = 1;
    return (int) (need < cap ? need : cap);
}

extern "C" {

const char * sq_last_error(void) { return sq_err; }
const char * sq_version(void) { return "shenqi_b200 0.1 (sm_100a)"; }

int sq_ctx_create(sq_ctx ** out, int device)
{
    if(!out) { sq_set_error("sq_ctx_create: null out pointer"); return SQ_ERR_ARG; }
    *out = nullptr;
    int ndev = 0;
    if(cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        sq_set_error("sq_ctx_create: no CUDA device - shenqi_b200 has no CPU path");
        return SQ_ERR_NO_DEVICE;
    }
    if(device < 0 || device >= ndev) { sq_set_error("sq_ctx_create: device %d of %d", device, ndev); return SQ_ERR_ARG; }
    cudaDeviceProp prop;
    SQ_CUDA(cudaGetDeviceProperties(&prop, device));
    if(prop.major < 10) {
        sq_set_error("sq_ctx_create: device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
        return SQ_ERR_NO_DEVICE;
    }
    SQ_CUDA(cudaSetDevice(device));
    sq_ctx * c = new sq_ctx();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    SQ_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for(int i = 0; i < 16; i++)
        SQ_CUDA(cudaEventCreate(&c->ev[i]));
    *out = c;
    return SQ_OK;
}

void sq_ctx_destroy(sq_ctx * c)
{
    if(!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    sq_pm_destroy(c);
    sq_dist_destroy(c);
    sq_comm_destroy(c);
    sq_buf * bufs[] = {&c->pos_in, &c->mass_in, &c->id_in, &c->type_in, &c->oldacc_in, &c->keys_in, &c->pos, &c->mass, &c->oldacc, &c->keys,
                       &c->perm, &c->rank, &c->sort_tmp_keys, &c->sort_tmp_vals, &c->sort_hist, &c->scan_tmp, &c->flag_tmp, &c->t_cm, &c->t_cl,
                       &c->t_lk, &c->t_parent, &c->t_rot, &c->t_lvl, &c->t_counters, &c->tree_keys, &c->tree_order, &c->tree_pos, &c->tree_mass,
                       &c->active_sorted, &c->w_accel, &c->w_pot, &c->w_stats, &c->w_targets, &c->w_table, &c->w_out_accel, &c->w_out_pot,
                       &c->w_counter, &c->w_icm, &c->w_icl, &c->w_ipart, &c->pm_real[0], &c->pm_real[1], &c->pm_real[2], &c->pm_real[3], &c->pm_cplx[0], &c->pm_cplx[1],
                       &c->pm_cplx[2], &c->pm_cplx[3], &c->pm_out, &c->slab_tab, &c->route_tmp, &c->oldacc_tmp, &c->stage_a, &c->stage_b, &c->act_tmp, &c->step_tmp,
                       &c->d_leafrange, &c->d_topnodes, &c->d_topleaves, &c->d_treenode, &c->t_top, &c->let_leafdata, &c->let_mask, &c->let_scan_n, &c->let_scan_p,
                       &c->let_roots, &c->let_send, &c->let_recv, &c->let_boxes, &c->w_gcost};
    for(sq_buf * b : bufs) b->release();
    for(int i = 0; i < 16; i++) if(c->ev[i]) cudaEventDestroy(c->ev[i]);
    for(int i = 0; i < 4; i++) if(c->evx[i]) cudaEventDestroy(c->evx[i]);
    if(c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if(c->stream && c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

int sq_set_params(sq_ctx * c, const sq_params * par)
{
    if(!c || !par) { sq_set_error("sq_set_params: null argument"); return SQ_ERR_ARG; }
    if(!(par->BoxSize > 0) || par->Nmesh <= 0 || !(par->G > 0)) { sq_set_error("sq_set_params: BoxSize, Nmesh and G must be positive"); return SQ_ERR_ARG; }
    if(par->WindowType != 1 && par->WindowType != 2) { sq_set_error("sq_set_params: ShortRangeForceWindowType must be 1 (exact) or 2 (erfc)"); return SQ_ERR_ARG; }
    // gravshort-tree2.cpp:42-46
    if(par->WindowType == 1 && par->Asmth != 1.5) {
        sq_set_error("The short range force window is calibrated for Asmth = 1.5, but running with %g", par->Asmth);
        return SQ_ERR_ARG;
    }
    c->par = *par;
    c->have_params = true;
    return SQ_OK;
}

int sq_set_walk_raw(sq_ctx * c, const sq_walk_raw * raw)
{
    if(!c) return SQ_ERR_ARG;
    if(!raw) { c->have_raw = false; return SQ_OK; }
    if(raw->ntab != SQ_NTAB || !raw->force_table || !raw->pot_table) { sq_set_error("sq_set_walk_raw: tables of %d entries expected (NGRAVTAB)", SQ_NTAB); return SQ_ERR_ARG; }
    if(!(raw->cellsize > 0) || !(raw->BoxSize > 0) || !(raw->G > 0) || !(raw->table_dx > 0)) { sq_set_error("sq_set_walk_raw: BoxSize, cellsize, G and table_dx must be positive"); return SQ_ERR_ARG; }
    c->raw = *raw;
    memcpy(c->raw_ft, raw->force_table, sizeof(float) * SQ_NTAB);
    memcpy(c->raw_pt, raw->pot_table, sizeof(float) * SQ_NTAB);
    c->raw.force_table = c->raw_ft; c->raw.pot_table = c->raw_pt;
    c->have_raw = true;
    c->raw_serial++;
    return SQ_OK;
}

int sq_sync(sq_ctx * c)
{
    if(!c) return SQ_ERR_ARG;
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    return SQ_OK;
}

void sq_default_aos_layout(sq_aos_layout * lay)
{
    lay->stride = 160; lay->off_pos = 0; lay->off_mass = 28; lay->off_flags = 36;
    lay->off_fulltreegravaccel = 64; lay->off_gravpm = 88; lay->off_id = 136; lay->off_potential = 152;
}

static int sq_reset_particles(sq_ctx * c, int64_t n)
{
    if(n < 0 || n >= ((int64_t) 1 << 31)) { sq_set_error("particle count %lld out of range (MaxPart < 2^31, partmanager.c:22)", (long long) n); return SQ_ERR_ARG; }
    c->n = n;
    c->sorted = false; c->tree_built = false; c->walk_done = false; c->pm_done = false;
    c->have_skip = false; c->n_valid = n;
    SQ_TRY(c->pos_in.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
    SQ_TRY(c->mass_in.reserve(sizeof(float) * (size_t) (n + 1)));
    SQ_TRY(c->oldacc_in.reserve(sizeof(double) * (size_t) (n + 1)));
    SQ_CUDA(cudaMemsetAsync(c->oldacc_in.p, 0, sizeof(double) * (size_t) (n + 1), c->stream));
    return SQ_OK;
}

int sq_upload_particles_soa(sq_ctx * c, const double * pos, const float * mass, const uint64_t * ids, const unsigned char * types, int64_t n)
{
    if(!c || (n > 0 && (!pos || !mass))) { sq_set_error("sq_upload_particles_soa: null argument"); return SQ_ERR_ARG; }
    (void) ids;
    if(types)
        for(int64_t i = 1; i < n; i++)
            if(types[i] != types[0]) { sq_set_error("sq_upload_particles_soa: mixed particle types are not supported yet (DM-only path)"); return SQ_ERR_ARG; }
    SQ_TRY(sq_reset_particles(c, n));
    SQ_CUDA(cudaMemcpyAsync(c->pos_in.p, pos, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->mass_in.p, mass, sizeof(float) * (size_t) n, cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    return SQ_OK;
}

int sq_upload_particles_aos(sq_ctx * c, const void * base, int64_t n, const sq_aos_layout * lay_in)
{
    if(!c || (n > 0 && !base)) { sq_set_error("sq_upload_particles_aos: null argument"); return SQ_ERR_ARG; }
    if(!c->have_params) { sq_set_error("sq_upload_particles_aos: call sq_set_params first (G is needed for OldAcc)"); return SQ_ERR_ARG; }
    sq_aos_layout lay;
    if(lay_in) lay = *lay_in; else sq_default_aos_layout(&lay);
    SQ_TRY(sq_reset_particles(c, n));
    // stage the records on the device, unpack there (the AoS is 160 B/particle; 64 B of it are read)
    if(lay.stride < 40 || lay.off_flags + 4 > lay.stride || lay.off_pos + 24 > lay.stride || lay.off_gravpm + 24 > lay.stride ||
       lay.off_fulltreegravaccel + 24 > lay.stride || lay.off_mass + 4 > lay.stride) {
        sq_set_error("sq_upload_particles_aos: field offsets do not fit the record stride %zu", lay.stride);
        return SQ_ERR_ARG;
    }
    sq_buf stage;
    const size_t chunk = 1 << 22;
    SQ_TRY(stage.reserve(lay.stride * (size_t) (n < (int64_t) chunk ? n + 1 : chunk)));
    SQ_TRY(c->skip_in.reserve((size_t) n + 64));
    int * d_nskip = (int *) (c->skip_in.as<unsigned char>() + (((size_t) n + 15) & ~(size_t) 15));
    SQ_CUDA(cudaMemsetAsync(d_nskip, 0, 4, c->stream));
    int rc = SQ_OK;
    for(int64_t i0 = 0; i0 < n; i0 += (int64_t) chunk) {
        const int64_t m = (n - i0) < (int64_t) chunk ? (n - i0) : (int64_t) chunk;
        cudaError_t e = cudaMemcpyAsync(stage.p, (const unsigned char *) base + (size_t) i0 * lay.stride, lay.stride * (size_t) m, cudaMemcpyHostToDevice, c->stream);
        if(e != cudaSuccess) { sq_set_error("sq_upload_particles_aos: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; break; }
        sq_unpack_aos_kernel<<<api_grid(c, m), 256, 0, c->stream>>>(stage.as<unsigned char>(), m, lay, c->par.G, c->pos_in.as<double>() + 3 * i0,
                c->mass_in.as<float>() + i0, c->oldacc_in.as<double>() + i0, c->skip_in.as<unsigned char>() + i0, d_nskip);
        c->kernel_launches++;
        e = cudaStreamSynchronize(c->stream);
        if(e != cudaSuccess) { sq_set_error("sq_upload_particles_aos: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; break; }
    }
    stage.release();
    if(rc == SQ_OK) {
        int h_nskip = 0;
        SQ_CUDA(cudaMemcpyAsync(&h_nskip, d_nskip, 4, cudaMemcpyDeviceToHost, c->stream));
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        c->have_skip = h_nskip > 0;
        c->n_valid = n - h_nskip;
    }
    return rc;
}

int sq_set_oldacc(sq_ctx * c, const double * oldacc)
{
    if(!c || !oldacc) { sq_set_error("sq_set_oldacc: null argument"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaMemcpyAsync(c->oldacc_in.p, oldacc, sizeof(double) * (size_t) c->n, cudaMemcpyDefault, c->stream));   // host or device pointer
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    if(c->sorted) SQ_TRY(sq_gather_run(c));
    return sq_sync(c);
}

int sq_set_old_accel(sq_ctx * c, const double * fulltree, const double * gravpm)
{
    if(!c) return SQ_ERR_ARG;
    if(!c->have_params) { sq_set_error("sq_set_old_accel: call sq_set_params first"); return SQ_ERR_ARG; }
    const int64_t n = c->n;
    // staging in two scratch buffers the context already owns (not in use before the next walk / download)
    sq_buf & a = c->w_out_accel, & b = c->oldacc_tmp;
    if(fulltree) { SQ_TRY(a.reserve(sizeof(double) * 3 * (size_t) (n + 1))); SQ_CUDA(cudaMemcpyAsync(a.p, fulltree, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice, c->stream)); }
    if(gravpm) { SQ_TRY(b.reserve(sizeof(double) * 3 * (size_t) (n + 1))); SQ_CUDA(cudaMemcpyAsync(b.p, gravpm, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice, c->stream)); }
    if(n > 0) {
        sq_oldacc_kernel<<<api_grid(c, n), 256, 0, c->stream>>>(fulltree ? a.as<double>() : nullptr, gravpm ? b.as<double>() : nullptr, n, c->par.G, c->oldacc_in.as<double>());
        c->kernel_launches++;
        SQ_CUDA(cudaGetLastError());
    }
    if(c->sorted) SQ_TRY(sq_gather_run(c));
    return sq_sync(c);
}

int sq_ph_sort(sq_ctx * c)
{
    if(!c || !c->have_params) { sq_set_error("sq_ph_sort: call sq_set_params first"); return SQ_ERR_ARG; }
    const int64_t n = c->n;
    c->sorted = false; c->tree_built = false;
    SQ_TRY(sq_keys_run(c));
    SQ_TRY(c->keys.reserve(sizeof(uint64_t) * (size_t) (n + 1)));
    SQ_TRY(c->perm.reserve(sizeof(uint32_t) * (size_t) (n + 1)));
    SQ_CUDA(cudaMemcpyAsync(c->keys.p, c->keys_in.p, sizeof(uint64_t) * (size_t) n, cudaMemcpyDeviceToDevice, c->stream));
    SQ_TRY(sq_iota(c, c->perm.as<uint32_t>(), n));
    SQ_CUDA(cudaEventRecord(c->ev[1], c->stream));
    SQ_TRY(sq_sort_pairs(c, c->keys.as<uint64_t>(), c->perm.as<uint32_t>(), n, 0, 64));
    SQ_CUDA(cudaEventRecord(c->ev[2], c->stream));
    SQ_TRY(sq_gather_run(c));
    SQ_CUDA(cudaEventRecord(c->ev[3], c->stream));
    int flags[4] = {0, 0, 0, 0};
    SQ_CUDA(cudaMemcpyAsync(flags, c->t_counters.p, sizeof(flags), cudaMemcpyDeviceToHost, c->stream));
    SQ_CUDA(cudaStreamSynchronize(c->stream));
    float ms = 0;
    if(n > 0) {
        cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]); c->tm.keys_ms = ms;
        cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]); c->tm.sort_ms = ms;
        cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]); c->tm.gather_ms = ms;
    }
    c->geo_mismatch = flags[0] & 1;
    if(flags[0] & 2) { sq_set_error("Zero mass particle (forcetree.cpp:808)"); return SQ_ERR_MASS; }
    c->sorted = true;
    return SQ_OK;
}

int sq_get_keys(sq_ctx * c, uint64_t * keys)
{
    if(!c || !keys || !c->keys_in.p) { sq_set_error("sq_get_keys: no keys"); return SQ_ERR_ARG; }
    SQ_CUDA(cudaMemcpyAsync(keys, c->keys_in.p, sizeof(uint64_t) * (size_t) c->n, cudaMemcpyDeviceToHost, c->stream));
    return sq_sync(c);
}

int sq_get_sorted(sq_ctx * c, uint64_t * sorted_keys, int32_t * perm)
{
    if(!c || !c->sorted) { sq_set_error("sq_get_sorted: call sq_ph_sort first"); return SQ_ERR_ARG; }
    if(sorted_keys) SQ_CUDA(cudaMemcpyAsync(sorted_keys, c->keys.p, sizeof(uint64_t) * (size_t) c->n, cudaMemcpyDeviceToHost, c->stream));
    if(perm) SQ_CUDA(cudaMemcpyAsync(perm, c->perm.p, sizeof(uint32_t) * (size_t) c->n, cudaMemcpyDeviceToHost, c->stream));
    return sq_sync(c);
}

int sq_tree_build(sq_ctx * c, const int32_t * active, int64_t nactive, sq_tree_info * info)
{
    if(!c) return SQ_ERR_ARG;
    if(active && nactive < 0) { sq_set_error("sq_tree_build: negative nactive"); return SQ_ERR_ARG; }
    return sq_tree_run(c, active, nactive, info);
}

int sq_tree_export_nodes(sq_ctx * c, void * nodes, int64_t firstnode, int upload_order)
{
    if(!c || !nodes) return SQ_ERR_ARG;
    return sq_tree_export(c, nodes, firstnode, upload_order, -1);
}

int sq_tree_export_nodes_ex(sq_ctx * c, void * nodes, int64_t firstnode, int64_t lastnode, int upload_order)
{
    if(!c || !nodes) return SQ_ERR_ARG;
    return sq_tree_export(c, nodes, firstnode, upload_order, lastnode);
}

int sq_tree_local_nodes(sq_ctx * c, int64_t * n)
{
    if(!c || !n || !c->tree_built) { sq_set_error("sq_tree_local_nodes: no tree"); return SQ_ERR_ARG; }
    *n = c->n_local_nodes;
    return SQ_OK;
}

int sq_tree_walk_device(sq_ctx * c, const int32_t * active, int64_t nactive, sq_walk_stats * stats)
{
    if(!c) return SQ_ERR_ARG;
    return sq_walk_run(c, active, nactive, stats);
}

int sq_download_accel(sq_ctx * c, double * accel, double * potential)
{
    if(!c || !c->walk_done) { sq_set_error("sq_download_accel: no walk results"); return SQ_ERR_ARG; }
    const int64_t n = c->n;
    // only the walked targets are written, as the reference's reduce does (gravshort2.hpp:136-146)
    const bool all = c->walk_all && c->n_valid == n;
    const uint32_t * targets = c->walk_all ? nullptr : c->w_targets.as<uint32_t>();
    const long long m = c->walk_ntargets;
    if(accel) {
        SQ_TRY(c->w_out_accel.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
        // a partial walk leaves the other particles' entries as the caller passed them: bring those in first
        if(!all) SQ_CUDA(cudaMemcpyAsync(c->w_out_accel.p, accel, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice, c->stream));
        if(m > 0) {
            sq_unpermute_kernel<3><<<api_grid(c, m), 256, 0, c->stream>>>(c->w_accel.as<double>(), 3, 0, c->perm.as<uint32_t>(), targets, m, c->w_out_accel.as<double>());
            c->kernel_launches++;
        }
        SQ_CUDA(cudaMemcpyAsync(accel, c->w_out_accel.p, sizeof(double) * 3 * (size_t) n, cudaMemcpyDeviceToHost, c->stream));
    }
    if(potential && c->tree_full) {
        SQ_TRY(c->w_out_pot.reserve(sizeof(double) * (size_t) (n + 1)));
        if(!all) SQ_CUDA(cudaMemcpyAsync(c->w_out_pot.p, potential, sizeof(double) * (size_t) n, cudaMemcpyHostToDevice, c->stream));
        if(m > 0) {
            sq_unpermute_kernel<1><<<api_grid(c, m), 256, 0, c->stream>>>(c->w_pot.as<double>(), 1, 0, c->perm.as<uint32_t>(), targets, m, c->w_out_pot.as<double>());
            c->kernel_launches++;
        }
        SQ_CUDA(cudaMemcpyAsync(potential, c->w_out_pot.p, sizeof(double) * (size_t) n, cudaMemcpyDeviceToHost, c->stream));
    }
    return sq_sync(c);
}

// results of the last walk into the caller's DEVICE arrays, upload order (only the walked targets are written)
int sq_download_accel_device(sq_ctx * c, double * accel, double * potential)
{
    if(!c || !c->walk_done) { sq_set_error("sq_download_accel_device: no walk results"); return SQ_ERR_ARG; }
    const uint32_t * targets = c->walk_all ? nullptr : c->w_targets.as<uint32_t>();
    const long long m = c->walk_ntargets;
    if(accel && m > 0) { sq_unpermute_kernel<3><<<api_grid(c, m), 256, 0, c->stream>>>(c->w_accel.as<double>(), 3, 0, c->perm.as<uint32_t>(), targets, m, accel); c->kernel_launches++; }
    if(potential && c->tree_full && m > 0) { sq_unpermute_kernel<1><<<api_grid(c, m), 256, 0, c->stream>>>(c->w_pot.as<double>(), 1, 0, c->perm.as<uint32_t>(), targets, m, potential); c->kernel_launches++; }
    SQ_CUDA(cudaGetLastError());
    return SQ_OK;
}

int sq_tree_walk(sq_ctx * c, const int32_t * active, int64_t nactive, double * accel, double * potential, sq_walk_stats * stats)
{
    if(!c) return SQ_ERR_ARG;
    SQ_TRY(sq_walk_run(c, active, nactive, stats));
    return sq_download_accel(c, accel, potential);
}

int sq_pm_force_device(sq_ctx * c)
{
    if(!c || !c->have_params) return SQ_ERR_ARG;
    return sq_pm_run(c, false);
}

int sq_pm_force(sq_ctx * c, double * gravpm, double * potential)
{
    if(!c || !c->have_params) return SQ_ERR_ARG;
    SQ_TRY(sq_pm_run(c, false));
    const int64_t n = c->n;
    if(gravpm) {
        SQ_TRY(c->w_out_accel.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
        if(n > 0) { sq_unpermute_kernel<3><<<api_grid(c, n), 256, 0, c->stream>>>(c->pm_out.as<double>(), 4, 1, c->perm.as<uint32_t>(), nullptr, n, c->w_out_accel.as<double>()); c->kernel_launches++; }
        SQ_CUDA(cudaMemcpyAsync(gravpm, c->w_out_accel.p, sizeof(double) * 3 * (size_t) n, cudaMemcpyDeviceToHost, c->stream));
    }
    if(potential) {
        SQ_TRY(c->w_out_pot.reserve(sizeof(double) * (size_t) (n + 1)));
        if(n > 0) { sq_unpermute_kernel<1><<<api_grid(c, n), 256, 0, c->stream>>>(c->pm_out.as<double>(), 4, 0, c->perm.as<uint32_t>(), nullptr, n, c->w_out_pot.as<double>()); c->kernel_launches++; }
        SQ_CUDA(cudaMemcpyAsync(potential, c->w_out_pot.p, sizeof(double) * (size_t) n, cudaMemcpyDeviceToHost, c->stream));
    }
    return sq_sync(c);
}

// One PM step of the reference's run loop in one call (run.cpp:538 gravpm_force, then run.cpp:556-563 force_tree_full +
// grav_short_tree) with HOST buffers, the copies overlapped with the kernels on a second stream:
//   copy stream:  H2D FullTreeGravAccel, GravPM (the |a_old| inputs)        | D2H PM accelerations + potential
//   main stream:  H2D pos, mass -> keys, sort, tree, PM -> unpermute PM out | |a_old|, gather -> walk -> unpermute -> D2H
// so that only the first upload and the last download are exposed.  Results equal the separate calls'.
static int sq_gravity_step_body(sq_ctx * c, const double * pos, const float * mass, int64_t n, const double * fulltree, const double * gravpm_old,
                                double * accel, double * potential, double * gravpm, double * pm_potential, sq_walk_stats * stats);

int sq_gravity_step(sq_ctx * c, const double * pos, const float * mass, int64_t n, const double * fulltree, const double * gravpm_old,
                    double * accel, double * potential, double * gravpm, double * pm_potential, sq_walk_stats * stats)
{
    if(!c || !c->have_params) { sq_set_error("sq_gravity_step: call sq_set_params first"); return SQ_ERR_ARG; }
    if(n > 0 && (!pos || !mass)) { sq_set_error("sq_gravity_step: null positions or masses"); return SQ_ERR_ARG; }
    if(!c->copy_stream) {
        SQ_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        for(int i = 0; i < 4; i++) SQ_CUDA(cudaEventCreateWithFlags(&c->evx[i], cudaEventDisableTiming));
    }
    const int rc = sq_gravity_step_body(c, pos, mass, n, fulltree, gravpm_old, accel, potential, gravpm, pm_potential, stats);
    if(rc != SQ_OK) {
        // the caller's buffers must not be in flight when the call returns, whatever went wrong (keep the first message)
        cudaStreamSynchronize(c->copy_stream);
        cudaStreamSynchronize(c->stream);
    }
    return rc;
}

static int sq_gravity_step_body(sq_ctx * c, const double * pos, const float * mass, int64_t n, const double * fulltree, const double * gravpm_old,
                                double * accel, double * potential, double * gravpm, double * pm_potential, sq_walk_stats * stats)
{
    SQ_TRY(sq_reset_particles(c, n));
    SQ_CUDA(cudaMemcpyAsync(c->pos_in.p, pos, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice, c->stream));
    SQ_CUDA(cudaMemcpyAsync(c->mass_in.p, mass, sizeof(float) * (size_t) n, cudaMemcpyHostToDevice, c->stream));
    const bool have_old = fulltree || gravpm_old;
    if(have_old) {
        // behind the positions on the bus: the sort can start as soon as those are in
        SQ_CUDA(cudaEventRecord(c->evx[3], c->stream));
        SQ_CUDA(cudaStreamWaitEvent(c->copy_stream, c->evx[3], 0));
        if(fulltree) { SQ_TRY(c->stage_a.reserve(sizeof(double) * 3 * (size_t) (n + 1))); SQ_CUDA(cudaMemcpyAsync(c->stage_a.p, fulltree, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice, c->copy_stream)); }
        if(gravpm_old) { SQ_TRY(c->stage_b.reserve(sizeof(double) * 3 * (size_t) (n + 1))); SQ_CUDA(cudaMemcpyAsync(c->stage_b.p, gravpm_old, sizeof(double) * 3 * (size_t) n, cudaMemcpyHostToDevice, c->copy_stream)); }
        SQ_CUDA(cudaEventRecord(c->evx[0], c->copy_stream));
    }
    SQ_TRY(sq_ph_sort(c));
    SQ_TRY(sq_tree_run(c, nullptr, 0, nullptr));
    const bool want_pm = gravpm || pm_potential;
    if(want_pm) {
        SQ_TRY(sq_pm_run(c, false));
        SQ_TRY(c->w_out_accel.reserve(sizeof(double) * 3 * (size_t) (n + 1)));
        SQ_TRY(c->w_out_pot.reserve(sizeof(double) * (size_t) (n + 1)));
        if(n > 0) {
            if(gravpm) { sq_unpermute_kernel<3><<<api_grid(c, n), 256, 0, c->stream>>>(c->pm_out.as<double>(), 4, 1, c->perm.as<uint32_t>(), nullptr, n, c->w_out_accel.as<double>()); c->kernel_launches++; }
            if(pm_potential) { sq_unpermute_kernel<1><<<api_grid(c, n), 256, 0, c->stream>>>(c->pm_out.as<double>(), 4, 0, c->perm.as<uint32_t>(), nullptr, n, c->w_out_pot.as<double>()); c->kernel_launches++; }
        }
        SQ_CUDA(cudaEventRecord(c->evx[1], c->stream));
        SQ_CUDA(cudaStreamWaitEvent(c->copy_stream, c->evx[1], 0));
        if(gravpm) SQ_CUDA(cudaMemcpyAsync(gravpm, c->w_out_accel.p, sizeof(double) * 3 * (size_t) n, cudaMemcpyDeviceToHost, c->copy_stream));
        if(pm_potential) SQ_CUDA(cudaMemcpyAsync(pm_potential, c->w_out_pot.p, sizeof(double) * (size_t) n, cudaMemcpyDeviceToHost, c->copy_stream));
        SQ_CUDA(cudaEventRecord(c->evx[2], c->copy_stream));
    }
    if(have_old) {
        SQ_CUDA(cudaStreamWaitEvent(c->stream, c->evx[0], 0));
        if(n > 0) {
            sq_oldacc_kernel<<<api_grid(c, n), 256, 0, c->stream>>>(fulltree ? c->stage_a.as<double>() : nullptr, gravpm_old ? c->stage_b.as<double>() : nullptr, n, c->par.G, c->oldacc_in.as<double>());
            c->kernel_launches++;
            SQ_CUDA(cudaGetLastError());
        }
        SQ_TRY(sq_gather_run(c));
    }
    SQ_TRY(sq_walk_run(c, nullptr, 0, stats));
    // the output staging buffers are shared with the PM download
    if(want_pm) SQ_CUDA(cudaStreamWaitEvent(c->stream, c->evx[2], 0));
    int rc = sq_download_accel(c, accel, potential);
    cudaError_t e = cudaStreamSynchronize(c->copy_stream);
    if(rc == SQ_OK && e != cudaSuccess) { sq_set_error("sq_gravity_step: %s", cudaGetErrorString(e)); rc = SQ_ERR_CUDA; }
    return rc;
}

int sq_pm_paint(sq_ctx * c)
{
    if(!c || !c->have_params) return SQ_ERR_ARG;
    return sq_pm_run(c, true);
}

int sq_pm_get_mesh(sq_ctx * c, double * mesh)
{
    if(!c || !mesh || !c->pm_real[0].p) { sq_set_error("sq_pm_get_mesh: no mesh"); return SQ_ERR_ARG; }
    const size_t N = (size_t) c->par.Nmesh;
    SQ_CUDA(cudaMemcpyAsync(mesh, c->pm_real[0].p, sizeof(double) * N * N * N, cudaMemcpyDeviceToHost, c->stream));
    return sq_sync(c);
}

int sq_get_timings(sq_ctx * c, sq_timings * t)
{
    if(!c || !t) return SQ_ERR_ARG;
    *t = c->tm;
    t->kernel_launches = c->kernel_launches;
    return SQ_OK;
}

int sq_reset_timings(sq_ctx * c)
{
    if(!c) return SQ_ERR_ARG;
    memset(&c->tm, 0, sizeof(c->tm));
    c->kernel_launches = 0;
    return SQ_OK;
}

} /* extern "C" */

// ------------------------------------------------------------------------------------ FP32 FMA peak probe
// Roofline denominator for the interaction kernel (MEASURED_PEAKS.json has HBM and bf16 only):
// 16 independent FFMA chains per thread, all SMs, 2 flop per FFMA.
__global__ void __launch_bounds__(256) sq_ffma_probe_kernel(float * out, int iters, float a, float b)
{
    float x[16];
    #pragma unroll
    for(int k = 0; k < 16; k++) x[k] = (float) (threadIdx.x + k);
    for(int it = 0; it < iters; it++) {
        #pragma unroll
        for(int k = 0; k < 16; k++) x[k] = fmaf(x[k], a, b);
    }
    float s = 0;
    #pragma unroll
    for(int k = 0; k < 16; k++) s += x[k];
    if(s == 123.456f) out[0] = s;
}

extern "C" int sq_measure_fp32_peak(sq_ctx * c, double * tflops)
{
    if(!c || !tflops) return SQ_ERR_ARG;
    sq_buf o;
    SQ_TRY(o.reserve(64));
    const int iters = 4096, blocks = c->num_sms * 8;
    double best = 0;
    for(int rep = 0; rep < 5; rep++) {
        SQ_CUDA(cudaEventRecord(c->ev[0], c->stream));
        sq_ffma_probe_kernel<<<blocks, 256, 0, c->stream>>>(o.as<float>(), iters, 1.0000001f, 1e-7f);
        SQ_CUDA(cudaEventRecord(c->ev[1], c->stream));
        SQ_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
        const double fl = 2.0 * 16.0 * iters * 256.0 * blocks;
        const double t = fl / (ms * 1e-3) / 1e12;
        if(rep > 0 && t > best) best = t;
    }
    o.release();
    *tflops = best;
    return SQ_OK;
}
