/* shenqi_b200 - B200-native (sm_100a) short-range TreePM gravity for MP-Gadget/shenqi.
 *
 * Thin C ABI underneath the reference's C++ entry points.  Plain pointers and
 * sizes only; no exceptions cross this boundary; every call returns SQ_OK or a negative error code
 * and sq_last_error() holds the message.  One context per process / GPU.  All calls are made from
 * one host thread; the library is multi-stream inside.  The caller keeps ownership of every host
 * array; the library owns its device pools.  There is NO CPU fallback: without a CUDA device
 * sq_ctx_create fails with SQ_ERR_NO_DEVICE.
 *
 * Reference interfaces each entry point stands behind (paths under MP-Gadget/shenqi):
 *   sq_set_params            set_gravshort_treepar / gravshort_set_softenings / gravpm_init_periodic
 *                            (libgadget/gravity.h:64-80, libgadget/gravshort-tree2.cpp:78-122)
 *   sq_upload_particles_aos  reads struct particle_data in place (libgadget/partmanager.h:9-71)
 *   sq_ph_sort               PEANO() + slots_gc_sorted (libgadget/utils/peano.h:15, slotsmanager.cpp:419)
 *   sq_tree_build            force_tree_full / force_tree_active_moments (libgadget/forcetree.h:127-140)
 *   sq_tree_export_nodes     the struct NODE image other modules walk (libgadget/forcetree.h:38-66)
 *   sq_tree_walk             grav_short_tree (libgadget/gravity.h:89, gravshort-tree2.cpp:141)
 *   sq_set_walk_raw          the GravTreeParams argument of grav_short_tree_cuda (libgadget/gravshort2.hpp:21-55, 440-443)
 *   sq_pm_force              gravpm_force -> petapm_force (libgadget/gravity.h:86, petapm.cpp:472)
 *   sq_drift_device          drift_all_particles (libgadget/drift.h:9, drift.cpp:81)
 *   sq_kick_device           apply_hierarchical_grav_kick (libgadget/timestep.cpp:249)
 *   sq_active_sublist_device build_active_sublist (libgadget/timestep.cpp:111, 1373)
 *   sq_gravity_timebins_device / sq_timebin_pushdown_device
 *                            the bin loop and push-down of hierarchical_gravity_accelerations (libgadget/timestep.cpp:351-407)
 *   sq_ctx_create_dist       one rank per GPU (README.rst:67): the MPI_COMM_WORLD of the tree path as an NCCL communicator
 *   sq_domain_set            struct DomainDecomp: topnode_data / topleaf_data / task_data (libgadget/domain.h:12-43)
 *   sq_domain_from_histogram domain_determine_global_toptree + domain_assign_topleaves_balanced (libgadget/domain.cpp:620-800, 1040-1160)
 *   sq_domain_assign_balanced domain_assign_topleaves_balanced, leaf for leaf (libgadget/domain.cpp:620-778)
 *   sq_tree_build_dist       force_tree_create_topnodes (forcetree.cpp:654-686, 868-933), force_exchange_pseudodata /
 *                            force_treeupdate_pseudos (:1155-1282), and - as a once-per-build import of remote sub-trees -
 *                            the export / ghost-walk / reduce cycle of TreeWalk::run_on_queue (treewalk2.h:282-373, 618-812)
 *   sq_toptree_export_count  GravTopTreeWalk::toptree_visit<TOPTREE_COUNT> (libgadget/gravshort2.hpp:362-438)
 *   sq_bigfile_*             petaio_read_block over depends/bigfile (libgadget/petaio.cpp:590-700, 908-1035)
 * The compiled reference-side bindings that call these live in shenqi_b200/host/ (INTEGRATION.md).
 */
#ifndef SHENQI_B200_H
#define SHENQI_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SQ_OK 0
#define SQ_ERR_NO_DEVICE (-1)     /* no CUDA device / wrong architecture: there is no CPU path */
#define SQ_ERR_CUDA (-2)          /* a CUDA runtime or cuFFT call failed */
#define SQ_ERR_ARG (-3)           /* bad argument or call order */
#define SQ_ERR_NODE_POOL (-4)     /* tree needs more nodes than the pool holds even after growth (TreeAllocFactor > 3, forcetree.cpp:215-231) */
#define SQ_ERR_TREE_DEPTH (-5)    /* more than 8 particles share one 2^-21 cell */
#define SQ_ERR_NONFINITE (-7)     /* a non-finite position after the drift (drift.cpp:70-73) */
#define SQ_ERR_MASS (-6)          /* a particle with Mass <= 0 (forcetree.cpp:808) */

typedef struct sq_ctx sq_ctx;

/* Knobs of the path: parameter-file names in comments (gadget/params.cpp:29-33, 84-104, 161-165). */
typedef struct sq_params {
    double BoxSize;            /* BoxSize */
    int Nmesh;                 /* Nmesh */
    double Asmth;              /* Asmth (1.5) */
    double G;                  /* Newton's constant, internal units */
    double ErrTolForceAcc;     /* ErrTolForceAcc (0.002) */
    double BHOpeningAngle;     /* BHOpeningAngle (0.175) */
    double MaxBHOpeningAngle;  /* MaxBHOpeningAngle (0.9) */
    int TreeUseBH;             /* TreeUseBH: 0 relative criterion, >0 Barnes-Hut; >1 = BH for the first call only */
    double TreeRcut;           /* TreeRcut (6), in units of Asmth*cellsize */
    double GravitySoftening;   /* comoving Plummer-equivalent softening length epsilon (FORCE_SOFTENING() = 2.8 epsilon) */
    int WindowType;            /* ShortRangeForceWindowType: 1 exact (table), 2 erfc */
    double rho0;               /* Omega0 * rho_crit, for the potential self term */
    double TreeAllocFactor;    /* node pool = factor * N (+ growth x1.15 up to 3.0); 0 -> 0.9 (run.cpp:257) */
} sq_params;

/* Byte offsets into the caller's array-of-structs particle record (defaults = struct particle_data). */
typedef struct sq_aos_layout {
    size_t stride;             /* 160 */
    size_t off_pos;            /* 0   double[3] */
    size_t off_mass;           /* 28  float */
    size_t off_flags;          /* 36  bit0 IsGarbage, bit1 Swallowed; byte +3 = Type */
    size_t off_fulltreegravaccel; /* 64 double[3] */
    size_t off_gravpm;         /* 88  double[3] */
    size_t off_id;             /* 136 uint64 */
    size_t off_potential;      /* 152 double */
} sq_aos_layout;

typedef struct sq_walk_stats {
    int64_t ntargets;          /* active particles walked */
    int64_t ngroups;           /* warps' worth of targets */
    int64_t interactions;      /* (target, source) pairs evaluated = the reference's ninteractions */
    int64_t list_entries;      /* shared interaction-list entries evaluated (x group width = lane slots issued) */
    int64_t nodes_tested;      /* (node, group) opening evaluations */
    int64_t pair_tests;        /* (target, node) pairs decided in FP32 with error intervals (tier 2) */
    int64_t exact_tests;       /* (target, node) pairs that needed the reference's FP64 expressions (tier 3) */
    double  kernel_ms;         /* device time of the walk kernel (CUDA events) */
    int64_t pseudo_opened;     /* (target, pseudo node) pairs that opened a pseudo node with nothing imported below it: must be 0 */
} sq_walk_stats;

typedef struct sq_tree_info {
    int64_t numnodes;          /* nodes created */
    int64_t numparticles;      /* particles in the tree */
    int64_t maxnodes;          /* pool size used */
    int     maxlevel;          /* deepest level (root = 0) */
    int     geo_mismatch;      /* 1 if some particle's geometric octants differ from its key digits (handled) */
} sq_tree_info;

/* The parameters of ONE grav_short_tree call exactly as the reference hands them to its GPU hook
 * grav_short_tree_cuda(): the fields of class GravTreeParams (libgadget/gravshort2.hpp:21-55) and the two
 * float tables of GravShortTable (libgadget/gravity.h:32-61).  When set, sq_tree_walk uses these instead of
 * deriving them from sq_params (cellsize = BoxSize/Nmesh, Rcut = TreeRcut*Asmth*cellsize, ...). */
typedef struct sq_walk_raw {
    double BoxSize;
    double cellsize;           /* BoxSize / Nmesh */
    double Rcut;               /* TreeRcut * Asmth * cellsize */
    double G;
    double cbrtrho0;
    double ForceSoftening;     /* 2.8 * epsilon */
    double ErrTolForceAcc;
    double BHOpeningAngle2;    /* already MaxBHOpeningAngle^2 when TreeUseBH == 0 */
    int TreeUseBH;
    int ntab;                  /* NGRAVTAB = 512 */
    const float * force_table; /* shortrange_table[ntab] */
    const float * pot_table;   /* shortrange_table_potential[ntab] */
    double table_dx;           /* table step in mesh cells (0.029354...) */
} sq_walk_raw;

const char * sq_last_error(void);
const char * sq_version(void);

int sq_ctx_create(sq_ctx ** ctx, int device);
void sq_ctx_destroy(sq_ctx * ctx);
int sq_set_params(sq_ctx * ctx, const sq_params * par);
int sq_sync(sq_ctx * ctx);
/* raw == NULL clears the override. */
int sq_set_walk_raw(sq_ctx * ctx, const sq_walk_raw * raw);

/* ---- particles ------------------------------------------------------------------------------- */
void sq_default_aos_layout(sq_aos_layout * lay);
/* Copy the fields the path needs out of the caller's AoS (host memory). */
int sq_upload_particles_aos(sq_ctx * ctx, const void * base, int64_t n, const sq_aos_layout * lay);
/* Same from separate host arrays; ids/types/oldacc may be NULL (ids = i+1, type 1, oldacc 0). */
int sq_upload_particles_soa(sq_ctx * ctx, const double * pos, const float * mass, const uint64_t * ids,
                            const unsigned char * types, int64_t n);
/* |FullTreeGravAccel + GravPM| / G per particle, in upload order (gravshort2.hpp:111-121); host arrays N x 3. */
int sq_set_old_accel(sq_ctx * ctx, const double * fulltreegravaccel, const double * gravpm);
/* Directly: oldacc[i] in upload order. */
int sq_set_oldacc(sq_ctx * ctx, const double * oldacc);

/* ---- Peano-Hilbert keys + sort (slots_gc_sorted) --------------------------------------------- */
/* Keys, stable radix sort by Key, gather into PH order on the device.  Collisionless particles of one type only (the
 * reference sorts by (TypeKey, Key), slotsmanager.cpp:395-412); garbage / swallowed particles uploaded through
 * sq_upload_particles_aos get the largest key and sort behind every valid particle (its TypeKey 255, :428-432). */
int sq_ph_sort(sq_ctx * ctx);
/* keys in upload order (before sort) */
int sq_get_keys(sq_ctx * ctx, uint64_t * keys);
/* sorted keys, and perm[i] = upload index of the particle at sorted position i */
int sq_get_sorted(sq_ctx * ctx, uint64_t * sorted_keys, int32_t * perm);

/* ---- force tree ------------------------------------------------------------------------------ */
/* Build the oct-tree (leaf capacity 8, root 1.001*Box) and monopole moments from the sorted
 * particles.  active: upload-order indices of the particles to put in the tree, or NULL for all
 * (force_tree_active_moments vs force_tree_full). */
/* `active` may be a host, managed (the reference's mymanagedmalloc lists, timestep.cpp:1319, 1376) or device pointer,
 * e.g. the list sq_active_sublist_device wrote: the copy is by unified virtual addressing. Same for sq_tree_walk*. */
int sq_tree_build(sq_ctx * ctx, const int32_t * active, int64_t nactive, sq_tree_info * info);
/* Write the tree as the reference's NODE[numnodes] image (120 B each), node numbers starting at
 * firstnode, particle numbers = sorted positions (or upload indices if upload_order != 0). */
int sq_tree_export_nodes(sq_ctx * ctx, void * nodes, int64_t firstnode, int upload_order);
/* Same with the caller's lastnode (= firstnode + maxnodes of its pool): the top leaves of other ranks are written as
 * PSEUDO nodes whose suns[0] = lastnode + leaf, as force_create_node_for_topnode leaves them (forcetree.cpp:903-906). */
int sq_tree_export_nodes_ex(sq_ctx * ctx, void * nodes, int64_t firstnode, int64_t lastnode, int upload_order);
/* nodes of this rank's own build (what the host image holds; sub-trees imported from other ranks come on top on the device) */
int sq_tree_local_nodes(sq_ctx * ctx, int64_t * n);

/* ---- short-range walk -------------------------------------------------------------------------- */
/* accel: N x 3 host doubles in upload order (already x G), written for active particles only.
 * potential: N host doubles or NULL; written only when the tree holds all particles
 * (full_particle_tree_flag), including the self terms of gravshort2.hpp:95-106.
 * TreeUseBH > 1 is reset to 0 after the call (gravshort-tree2.cpp:170-171). */
int sq_tree_walk(sq_ctx * ctx, const int32_t * active, int64_t nactive, double * accel, double * potential,
                 sq_walk_stats * stats);
/* Same, results left on the device (benchmark / chained use); fetch with sq_download_accel. */
int sq_tree_walk_device(sq_ctx * ctx, const int32_t * active, int64_t nactive, sq_walk_stats * stats);
int sq_download_accel(sq_ctx * ctx, double * accel, double * potential);
/* Same into DEVICE arrays (upload order), queued on the context's stream, no host synchronisation. */
int sq_download_accel_device(sq_ctx * ctx, double * accel, double * potential);
/* Work estimate per particle of the last walk, upload order: interactions of its target group / group width (what the
 * reference keeps as GravCost for the domain balance, domain.cpp:452-527).  cost: host or device float[N]. */
int sq_walk_particle_cost(sq_ctx * ctx, float * cost);

/* ---- particle mesh --------------------------------------------------------------------------- */
/* CIC paint -> r2c FFT -> potential + 3 force transfer functions -> 4 c2r FFTs -> fused CIC readout.
 * gravpm: N x 3 host doubles (upload order); potential: N doubles or NULL (PM part only, added to nothing). */
int sq_pm_force(sq_ctx * ctx, double * gravpm, double * potential);
int sq_pm_force_device(sq_ctx * ctx);
/* stages, for tests: density mesh Nmesh^3 (z fastest) after paint */
int sq_pm_paint(sq_ctx * ctx);
int sq_pm_get_mesh(sq_ctx * ctx, double * mesh);

/* ---- one PM step in one call ------------------------------------------------------------------ */
/* What run() does on a PM step for gravity (libgadget/run.cpp:538 gravpm_force; run.cpp:556-563 force_tree_full +
 * grav_short_tree): upload, PH sort, tree, PM force, short-range walk, download - with HOST buffers (pinned for overlap),
 * the |a_old| inputs uploaded and the PM results downloaded on a second stream underneath the kernels.
 * pos N x 3, mass N; fulltree / gravpm_old: N x 3 each or NULL (Part[i].FullTreeGravAccel, Part[i].GravPM of the
 * previous step, gravshort2.hpp:111-121); accel N x 3 (x G), potential N or NULL (tree part);
 * gravpm N x 3 and pm_potential N: PM outputs, both NULL = no PM force this step.  Results equal the separate calls'. */
int sq_gravity_step(sq_ctx * ctx, const double * pos, const float * mass, int64_t n, const double * fulltree, const double * gravpm_old,
                    double * accel, double * potential, double * gravpm, double * pm_potential, sq_walk_stats * stats);

/* ---- multi-GPU: one rank per GPU (README.rst:67 of the reference), exchanges issued by the host over NCCL ------
 * Everything below takes DEVICE pointers and runs on the context's stream; the host layer (shenqi_b200/dist.py with
 * torch.distributed, or a C++ host with raw NCCL) issues the collectives between the calls.
 * Tree side (stands where the reference exchanges top-leaf moments and exports/imports walks,
 * libgadget/forcetree.cpp:1144-1282, treewalk2.h:282-373, 742-812): every rank installs the all-gathered, globally
 * PH-sorted particle set, builds the whole tree and walks only its own range of targets.
 * PM side (stands behind petapm_force, libgadget/petapm.cpp:371-472, with the region/pencil exchange of
 * petapm.cpp:692-1107 replaced by routing particles to x-slab owners and the heFFTe transposes, petapm.cpp:281-303,
 * by an all-to-all between the 2-D and 1-D cuFFT passes). */
#define SQ_PTR_SORTED_KEYS 0      /* uint64[n]   */
#define SQ_PTR_SORTED_POS 1       /* double[n][3] */
#define SQ_PTR_SORTED_MASS 2      /* float[n]    */
#define SQ_PTR_SORTED_PERM 3      /* uint32[n]: upload index of the particle at sorted position i */
#define SQ_PTR_WALK_ACCEL 4       /* double[n][3], by sorted position, already x G */
#define SQ_PTR_WALK_POT 5         /* double[n] */
#define SQ_PTR_PM_OUT 6           /* double[n][4]: PM potential, fx, fy, fz by sorted position */
/* Run every later call of this context on the caller's CUDA stream (e.g. torch's current stream). */
int sq_ctx_set_stream(sq_ctx * ctx, void * cuda_stream);
int sq_device_ptr(sq_ctx * ctx, int which, void ** ptr, int64_t * count);
/* sq_upload_particles_soa from device arrays (pos: n x 3 doubles, mass: n floats) */
int sq_upload_particles_device(sq_ctx * ctx, const double * pos, const float * mass, int64_t n);
/* 1 if sq_ph_sort found particles whose geometric octants differ from their key digits (exactly on a cell face) */
int sq_get_geo_mismatch(sq_ctx * ctx, int * flag);
int sq_set_sorted_device(sq_ctx * ctx, const uint64_t * keys, const double * pos, const float * mass, int64_t n, int geo_mismatch);
int sq_set_oldacc_range_device(sq_ctx * ctx, int64_t first, int64_t count, const double * oldacc);
int sq_tree_walk_range(sq_ctx * ctx, int64_t first, int64_t count, sq_walk_stats * stats);
int sq_pm_slab_route(sq_ctx * ctx, const double * pos, const float * mass, int64_t n, int nranks, double * send4, int32_t * slot, int64_t * counts);
int sq_pm_slab_unroute(sq_ctx * ctx, const double * back4, const int32_t * slot, int64_t n, double * out4);
int sq_pm_slab_paint(sq_ctx * ctx, const double * part4, int64_t m, int rank, int nranks, double * slab);
int sq_pm_slab_fft_fwd(sq_ctx * ctx, double * slab, int nranks, double * work, double * packed);
int sq_pm_slab_fft_x(sq_ctx * ctx, double * data, int nranks, int inverse);
int sq_pm_slab_transfer(sq_ctx * ctx, const double * rho, int rank, int nranks, double * out4);
/* phi_k only; the forces are then differenced in real space from the potential slab and two / three planes of each
 * x-neighbour (4-point stencil = the reference's k-space filter of gravpm.cpp:447-475 in real space). */
int sq_pm_slab_transfer_phi(sq_ctx * ctx, const double * rho, int rank, int nranks, double * phik);
int sq_pm_slab_gradient(sq_ctx * ctx, double * phi, const double * lo2, const double * hi3, int nranks, double * ext, double * fx, double * fy, double * fz);
int sq_pm_slab_fft_inv(sq_ctx * ctx, const double * recv, int nranks, double * work, double * slab);
int sq_pm_slab_readout(sq_ctx * ctx, const double * part4, int64_t m, int rank, int nranks, const double * s0, const double * s1,
                       const double * s2, const double * s3, double * out4);

/* ---- multi-rank tree: the reference's own partitioning (rows a5 top nodes, a7, a14, a15) ----------------
 * One rank per GPU.  The domain is the reference's DomainDecomp (libgadget/domain.h:12-43): a top tree over
 * Peano-Hilbert key ranges whose leaves are assigned to ranks.  Every rank builds the SAME top-level nodes
 * (force_tree_create_topnodes / force_create_node_for_topnode, forcetree.cpp:654-686, 868-933), the sub-trees of its own
 * top leaves, and pseudo nodes for the others; top-leaf moments are exchanged (force_exchange_pseudodata :1155-1206)
 * and the internal top nodes recomputed (force_treeupdate_pseudos :1212-1282).  Boundary walks: where the reference
 * exports a particle to the owner of every pseudo node it opens (GravTopTreeWalk::toptree_visit gravshort2.hpp:362-438,
 * export table treewalk2.h:618-677, query / result exchange :742-812, ghost walk :687-733), this library imports, once
 * per tree build, the part of each remote sub-tree that a local particle can reach - every node whose keep box
 * (|centre - x| <= Rcut + len/2 on every axis, gravshort2.hpp:152-167) meets the bounding box of a local top leaf,
 * with its true moments, and the particles of such leaves - and hangs it below the pseudo node.  The local walk then
 * makes, for every target, exactly the decisions the reference's primary + ghost walks make together: the interaction
 * set is the single-rank one (sq_walk_stats.interactions summed over ranks equals the one-rank count).
 * All collectives are raw NCCL on the context's stream, behind this ABI (ncclAllReduce for moments / boxes / counts,
 * grouped ncclSend / ncclRecv for the imports). */
#define SQ_NCCL_UNIQUE_ID_BYTES 128
/* same bytes as struct topnode_data / topleaf_data / task_data (libgadget/domain.h:12-31) */
typedef struct sq_topnode { uint64_t StartKey; int32_t Daughter; int32_t Shift; int32_t Leaf; int32_t pad_; } sq_topnode;   /* 24 B */
typedef struct sq_topleaf { int32_t Task; int32_t topnode; int32_t treenode; } sq_topleaf;                                  /* 12 B */
typedef struct sq_task { int32_t StartLeaf; int32_t EndLeaf; } sq_task;                                                     /*  8 B */

/* ncclGetUniqueId: call on one rank, hand the 128 bytes to every rank (MPI_Bcast in the reference's host). */
int sq_comm_unique_id(void * id128);
/* SURVEY 8b's sq_ctx_create(device, ncclUniqueId*, rank, nranks): a context that is one rank of nranks (collective). */
int sq_ctx_create_dist(sq_ctx ** ctx, int device, const void * nccl_unique_id, int rank, int nranks);
int sq_comm_rank(sq_ctx * ctx, int * rank, int * nranks);
/* TEST TRANSPORT: an id for sq_ctx_create_dist that makes the ranks - host threads of ONE process, each with its own
 * context, on one GPU or several - meet in an in-process hub instead of NCCL (device-to-device copies).  It lets the
 * multi-rank path run where NCCL cannot (two ranks on one GPU); production runs use sq_comm_unique_id. */
int sq_comm_loopback_id(void * id128, int key);
/* Install the domain for the following tree builds (arrays are copied; treenode is filled by the build and can be read
 * back with sq_domain_get_treenodes).  NULL / 0 removes it (trivial single-leaf domain).  Works with one rank too:
 * the tree then carries the forced top-level nodes of the domain, as the reference's does. */
int sq_domain_set(sq_ctx * ctx, const sq_topnode * topnodes, int ntopnodes, const sq_topleaf * topleaves, int ntopleaves,
                  const sq_task * tasks, int ntasks);
/* Pure host code: the consistency checks sq_domain_set applies to a DomainDecomp (daughters in key order with
 * StartKey / Shift as domain.cpp:849-868, every leaf numbered once, tasks own contiguous leaf ranges); *maxlevel = depth
 * of the deepest top node.  SQ_OK if the domain would be accepted. */
int sq_domain_validate(const sq_topnode * topnodes, int ntopnodes, const sq_topleaf * topleaves, int ntopleaves,
                       const sq_task * tasks, int ntasks, int * maxlevel);
int sq_domain_get_treenodes(sq_ctx * ctx, int32_t * treenode /* [ntopleaves], node numbers relative to the first node */);
/* Pure host code, no CUDA device needed: a top tree from a histogram of the particles' Peano-Hilbert keys over the
 * 8^level cells of level `level` (cell = key >> (63 - 3 level)), refined until no leaf holds more than
 * total / (overdecomposition * nranks) (domain.cpp:399, 1040-1160), leaves in key order cut into nranks contiguous
 * segments of equal cost (domain_assign_topleaves_balanced, domain.cpp:620-800); cost may be NULL (= count).
 * Arrays must hold max_topnodes entries; returns SQ_OK and the sizes. */
int sq_domain_from_histogram(const int64_t * count, const double * cost, int level, int nranks, int overdecomposition,
                             sq_topnode * topnodes, sq_topleaf * topleaves, sq_task * tasks, int max_topnodes,
                             int * ntopnodes, int * ntopleaves);
/* Pure host code: the reference's own assignment of top leaves to tasks (domain_assign_topleaves_balanced,
 * domain.cpp:620-778), leaf for leaf: cost[k] = integer cost of the k-th top leaf in Peano-Hilbert key order,
 * task[k] = its task.  With nsegment_per_task = 1 (what the reference calls it with: domain_balance,
 * domain.cpp:506-526) the tasks are contiguous in key order.  Fails if there are fewer leaves than
 * nranks * nsegment_per_task. */
int sq_domain_assign_balanced(const int64_t * cost, int ntopleaves, int nranks, int nsegment_per_task, int32_t * task);
/* Histogram of this rank's uploaded particles over the level-`level` cells, summed over ranks (ncclAllReduce when the
 * context is distributed); weights = per-particle cost in upload order (device or host pointer) or NULL. */
int sq_domain_histogram(sq_ctx * ctx, int level, int64_t * count, double * cost, const float * weights);
/* PEANO() keys (libgadget/utils/peano.h:15-21) of n positions that are already on the device; keys: device uint64[n]. */
int sq_keys_device(sq_ctx * ctx, const double * pos, int64_t n, uint64_t * keys);
/* Rank that owns each uploaded particle under the installed domain (domain_get_topleaf, domain.h:68-75); host array. */
int sq_domain_owner(sq_ctx * ctx, int32_t * owner);
/* Collective.  sq_tree_build under the installed domain: local sub-trees + top nodes, moment exchange, import of the
 * reachable parts of the remote sub-trees.  After it sq_tree_walk* walk the local particles against the global tree. */
typedef struct sq_let_info {
    int64_t local_nodes, local_particles;        /* this rank's tree before the import */
    int64_t imported_nodes, imported_particles;  /* received from the other ranks */
    int64_t exported_nodes, exported_particles;  /* sent to the other ranks (sum over peers) */
    int64_t bytes_sent, bytes_received;
    int64_t toptree_exports;                     /* (particle, rank) pairs the reference's toptree walk would export, if counted */
    double build_ms, moments_exchange_ms, let_mark_ms, let_exchange_ms, let_attach_ms;
} sq_let_info;
int sq_tree_build_dist(sq_ctx * ctx, sq_tree_info * info, sq_let_info * let);
/* GravTopTreeWalk::toptree_visit<TOPTREE_COUNT> (gravshort2.hpp:362-438) for every local particle over the top tree with
 * exchanged moments: the number of (particle, rank) exports the reference would make - the measured export fraction that
 * decides between query export and sub-tree import.  Host output. */
int sq_toptree_export_count(sq_ctx * ctx, int64_t * nexports, int64_t * nparticles_exporting);

/* ---- drift, hierarchical gravity kick and active sub-list on DEVICE arrays --------
 * All pointers except random_shift and nsub are device pointers on the context's device; work is queued on the
 * context's stream.  `skip` (may be NULL) holds one byte per particle, non-zero = IsGarbage || Swallowed.
 *
 * sq_drift_device replaces drift_all_particles / real_drift_particle (libgadget/drift.cpp:16-78, 81-99) for
 * collisionless particles: Pos += Vel * ddrift + random_shift, wrapped into (0, BoxSize]; skipped particles only
 * follow the shift.  A non-finite result is fatal in the reference (endrun 5): SQ_ERR_NONFINITE here.
 * sq_kick_device replaces the loop of apply_hierarchical_grav_kick (libgadget/timestep.cpp:270-287, 962-968):
 * Vel[pa] += Accel[pa] * gravkick over the active list (NULL = all of [0, nactive)).
 * sq_active_sublist_device replaces build_active_sublist (libgadget/timestep.cpp:1353-1397): the stable selection of
 * the entries of `active` (NULL = 0 .. nactive-1) with TimeBinGravity <= maxtimebin whose bin is active at
 * Ti_Current (is_timebin_active, timestep.cpp:132-139); *nsub receives the count (the call synchronises). */
int sq_drift_device(sq_ctx * ctx, double * pos, const double * vel, const unsigned char * skip, int64_t n, double ddrift,
                    const double * random_shift, double BoxSize);
int sq_kick_device(sq_ctx * ctx, double * vel, const double * accel, const int32_t * active, const unsigned char * skip,
                   int64_t nactive, double gravkick);
int sq_active_sublist_device(sq_ctx * ctx, const int32_t * active, int64_t nactive, const unsigned char * timebin,
                             const unsigned char * skip, int maxtimebin, int64_t Ti_Current, int32_t * sublist, int64_t * nsub);

/* Gravity time-bin assignment of hierarchical_gravity_accelerations (libgadget/timestep.cpp:351-377), per active particle:
 * get_timestep_gravity_dloga (:1029-1039, grav_acceleration2 :1012-1027) -> convert_timestep_to_ti (:157-174, with
 * TimeBinMgr::dti_from_dloga / ti_from_loga_snap, libgadget/timebinmgr.h:123-156) -> round_down_power_of_two
 * (libgadget/timebinmgr.cpp:190) -> get_timestep_bin (:1236-1250) -> min(bin, largest_active) -> TimeBinGravity.
 * The host fills the scalars from its Cosmology / TimeBinMgr state. */
typedef struct sq_timebin_params {
    double atime, hubble;              /* scale factor and H(a) at Ti_Current */
    double ErrTolIntAccuracy;          /* TimestepParams, timestep.cpp:58 */
    double MinSizeTimestep;            /* timestep.cpp:62 */
    double softening;                  /* FORCE_SOFTENING() / 2.8 */
    int64_t Ti_Current, dti_max;
    double loga;                       /* TimeBinMgr::loga_from_ti(Ti_Current) */
    int64_t lastsnap;                  /* Ti_Current >> TIMEBINS, clamped as timebinmgr.h:139-150 */
    int can_advance;                   /* lastsnap < NSyncPoints - 2 */
    double sync_loga[3];               /* SyncPoints[lastsnap .. lastsnap+2].loga (the third only if can_advance) */
    int largest_active;
} sq_timebin_params;
#define SQ_TIMEBINS 46                 /* libgadget/timebinmgr.h:16 */
/* accel = the acceleration store (or FullTreeGravAccel), gravpm = GravPM, both device double[n][3]; timebin = device
 * uint8[n], written for every active, non-skipped particle; counts[SQ_TIMEBINS + 1] (host) receives timebincounts of
 * this rank (the caller all-reduces them, timestep.cpp:378); *nbad = particles print_bad_timebin would report. */
int sq_gravity_timebins_device(sq_ctx * ctx, const sq_timebin_params * par, const double * accel, const double * gravpm,
                               const int32_t * active, const unsigned char * skip, int64_t nactive, unsigned char * timebin,
                               int64_t * counts, int64_t * nbad);
/* timestep.cpp:399-407: TimeBinGravity = min(TimeBinGravity, push_down_bin) over the active list. */
int sq_timebin_pushdown_device(sq_ctx * ctx, const int32_t * active, int64_t nactive, unsigned char * timebin, int push_down_bin);

/* ---- bigfile snapshot / IC reader (host code; petaio.cpp:590-700, 908-1035 over depends/bigfile) ----------
 * <file>/<block>/header ("DTYPE: .. NMEMB: .. NFILE: .." + one "%06X: rows : checksum : sysvsum" line per physical file),
 * raw little-endian rows in <file>/<block>/%06X, attributes in <file>/<block>/attr-v2.  Checksums (byte sum mod 2^32) are
 * verified for every physical file read in full. */
int sq_bigfile_block_info(const char * path, const char * block, char * dtype8, int * nmemb, int64_t * nrows, int * nfile);
/* rows [first, first + count) converted to `want` = "f8" | "f4" | "u8" | "i8" | "u4" | "i4" | "u1"; out: count x nmemb values */
int sq_bigfile_read_block(const char * path, const char * block, int64_t first, int64_t count, void * out, const char * want);
/* a numeric attribute as doubles: at most maxn values written, *n = its length */
int sq_bigfile_read_attr(const char * path, const char * block, const char * name, double * values, int maxn, int * n);
/* "<ptype>/Position" + "<ptype>/Mass" (or Header.MassTable[ptype] when non-zero) of rows [first, first + count) (count < 0:
 * to the end) into the context (sq_upload_particles_soa); *boxsize = Header.BoxSize if wanted */
int sq_upload_particles_bigfile(sq_ctx * ctx, const char * path, int ptype, int64_t first, int64_t count, double * boxsize);

/* ---- timing of the last device stages (CUDA events on the launching stream), milliseconds ------ */
typedef struct sq_timings {
    double keys_ms, sort_ms, gather_ms, tree_build_ms, tree_moments_ms, walk_ms;
    double pm_paint_ms, pm_fft_ms, pm_transfer_ms, pm_readout_ms;
    int64_t kernel_launches;   /* kernels of this library launched since sq_reset_timings */
} sq_timings;
int sq_get_timings(sq_ctx * ctx, sq_timings * t);
int sq_reset_timings(sq_ctx * ctx);
/* FFMA-pipe peak of this device in TFLOP/s (2 flop per FMA), measured: roofline denominator of the walk. */
int sq_measure_fp32_peak(sq_ctx * ctx, double * tflops);

#ifdef __cplusplus
}
#endif
#endif
