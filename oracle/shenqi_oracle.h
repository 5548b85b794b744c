/* TEST INFRASTRUCTURE ONLY - CPU restatement ("port") of the reference's short-range TreePM path.
 *
 * Plain C, double precision, one function per reference function, each citing the file:line it
 * follows (paths relative to /root/reference).  Pinned (tests/test_oracle_*.py) against
 *   - the 64 golden Peano-Hilbert keys of libgadget/tests/test_peano.cpp:102 (+ Hilbert-curve adjacency),
 *   - the reference's own code compiled into oracle/_ref/libshenqi_ref.so (keys, sort order, tree
 *     topology + moments, grav_short_tree accelerations, interaction counts, grav_short_pair),
 *   - the tree invariants of libgadget/tests/test_forcetree.cpp:21-169.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
 * The product (shenqi_b200/) never links or calls it.
 */
#ifndef SHENQI_ORACLE_H
#define SHENQI_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Same byte layout as the reference's struct NODE (libgadget/forcetree.h:29-66), 120 bytes. */
typedef struct sqo_node {
    int sibling;
    int father;
    double len;
    double center[3];
    double cofm[3];
    double mass;
    double hmax;
    int suns[8];
    int noccupied;
    unsigned int flags; /* bit0 InternalTopLevel, bit1 TopLevel, bit2 DependsOnLocalMass, bits3-4 ChildType */
} sqo_node;

#define SQO_PARTICLE_NODE 0
#define SQO_NODE_NODE 1
#define SQO_PSEUDO_NODE 2
#define SQO_CHILDTYPE(n) (((n)->flags >> 3) & 3u)
#define SQO_TOPLEVEL(n) (((n)->flags >> 1) & 1u)

/* Parameters of one grav_short_tree call: GravTreeParams, libgadget/gravshort2.hpp:21-55 */
typedef struct sqo_grav_params {
    double BoxSize;
    int Nmesh;
    double Asmth;
    double G;
    double ErrTolForceAcc;
    double BHOpeningAngle;
    double MaxBHOpeningAngle;
    int TreeUseBH;
    double TreeRcut;           /* in units of Asmth * cellsize */
    double GravitySoftening;   /* comoving Plummer-equivalent epsilon; ForceSoftening = 2.8 * epsilon */
    int WindowType;            /* 1 exact table, 2 erfc */
    double rho0;
    int update_potential;      /* tree->full_particle_tree_flag */
} sqo_grav_params;

/* libgadget/utils/peano.cpp:144-183 */
uint64_t sqo_peano_hilbert_key(int x, int y, int z, int bits);
/* libgadget/utils/peano.h:15-21 */
void sqo_peano_keys(const double * pos, int64_t n, double box, uint64_t * keys);
/* libgadget/slotsmanager.cpp:395-437: order by (TypeKey, Key); ties broken by input index */
void sqo_sort_order(const uint64_t * keys, const unsigned char * types, int64_t n, int64_t * perm);

/* libgadget/forcetree.cpp:653-686, 726-859, 480-519, 369-476, 1016-1142 with the single-rank
 * trivial domain of libgadget/tests/test_forcetree.cpp:294-314.  nodes[0] is the root
 * (node number firstnode).  Returns the number of nodes used, or -1 if maxnodes is too small. */
int64_t sqo_tree_build(const double * pos, const float * mass, int64_t n, double box,
                       sqo_node * nodes, int64_t maxnodes, int64_t firstnode);

/* Same, with the reference's "start from the last node touched if the particle is inside it" shortcut
 * (forcetree.cpp:811, an INCLUSIVE |2 dx| <= len test) switchable.  With the shortcut a particle lying
 * exactly on a cell face lands in whichever neighbour the previous particle went to, so the reference's
 * topology for such inputs (lattices) depends on insertion order and thread count; without it every
 * particle descends from the root by `Pos > centre` alone, which is the canonical placement. */
int64_t sqo_tree_build_opt(const double * pos, const float * mass, int64_t n, double box,
                           sqo_node * nodes, int64_t maxnodes, int64_t firstnode, int last_node_shortcut);

/* libgadget/gravshort2.hpp:227-358 (visit<PRIMARY>, apply_accn), :88-107 (postprocess), :111-121.
 * oldacc[i] = |FullTreeGravAccel+GravPM|/G.  accel (N x 3) is multiplied by G; potential may be NULL.
 * counts[i] = number of apply_accn calls (may be NULL).  Returns total interactions. */
int64_t sqo_grav_short_tree(const sqo_node * nodes, int64_t firstnode, const double * pos, const float * mass,
                            const double * oldacc, const int * active, int64_t nactive,
                            const sqo_grav_params * par, double * accel, double * potential, int64_t * counts);

/* libgadget/gravshort-pair.c:22-127: direct sum over r <= Rcut*Asmth*cellsize, O(N^2). accel x G. */
void sqo_grav_short_pair(const double * pos, const float * mass, int64_t n, const sqo_grav_params * par,
                         double * accel, double * potential);

/* libgadget/tests/test_gravity.cpp:44-83, 133-153: the O(N^2) periodic (27 images) checker of the reference's gravity test */
void sqo_direct_sum_27(const double * pos, const float * mass, int64_t n, double box, double G, double h, double * accn);

/* libgadget/petapm.cpp:1132-1183 + 1304-1310 on one periodic slab: mesh[Nmesh^3], z fastest. */
void sqo_cic_paint(const double * pos, const float * mass, int64_t n, double box, int nmesh, double * mesh);
/* libgadget/petapm.cpp:1132-1183 + gravpm.cpp:489-500: out[i] += sum_w mesh */
void sqo_cic_readout(const double * mesh, const double * pos, int64_t n, double box, int nmesh, double * out);
/* libgadget/gravpm.cpp:378-444 (potential_transfer, no neutrinos) and :448-487 (force_transfer), on a
 * numpy rfftn layout: rho_k[x][y][kz], kz = 0..Nmesh/2, complex interleaved.  out arrays same shape. */
void sqo_pm_transfer(const double * rho_k, int nmesh, double box, double asmth, double G,
                     double * pot_k, double * fx_k, double * fy_k, double * fz_k);

/* Window table as GravShortTable builds it (libgadget/gravshort-tree2.cpp:29-69): floats. */
void sqo_window_table(int windowtype, double asmth, float * force_tab, float * pot_tab, double * dx);

/* ---- drift, hierarchical gravity kick, active sub-list ---- */
/* libgadget/drift.cpp:16-81 (collisionless particles; skip[i] != 0 = IsGarbage || Swallowed) applied to all n
 * particles as drift_all_particles does (:84-103).  Returns the number of non-finite coordinates (fatal there). */
int64_t sqo_drift(double * pos, const double * vel, const unsigned char * skip, int64_t n, double ddrift,
                  const double * random_shift, double box);
/* libgadget/timestep.cpp:270-279 + 962-968: Vel[pa] += Accel[pa] * gravkick over the active list (NULL = all). */
void sqo_kick(double * vel, const double * accel, const int32_t * active, const unsigned char * skip, int64_t nactive, double gravkick);
/* libgadget/timestep.cpp:132-139 */
int sqo_is_timebin_active(int bin, int64_t current);
/* libgadget/timestep.cpp:1353-1397 (SubActivePredicate + build_active_sublist, std::copy_if = stable). Returns the count. */
int64_t sqo_active_sublist(const int32_t * active, int64_t nactive, const unsigned char * timebin, const unsigned char * skip,
                           int maxtimebin, int64_t ti_current, int32_t * sublist);

/* libgadget/timestep.cpp:351-377 (gravity time-bin assignment of hierarchical_gravity_accelerations) with
 * grav_acceleration2 / get_timestep_gravity_dloga (:1012-1039), convert_timestep_to_ti (:157-174),
 * TimeBinMgr::dti_from_dloga / ti_from_loga_snap (libgadget/timebinmgr.h:123-156), round_down_power_of_two
 * (libgadget/timebinmgr.cpp:189-202), get_timestep_bin (:1236-1250).  Same scalars as sq_timebin_params. */
typedef struct sqo_timebin_params {
    double atime, hubble, ErrTolIntAccuracy, MinSizeTimestep, softening;
    int64_t Ti_Current, dti_max;
    double loga;
    int64_t lastsnap;
    int can_advance;
    double sync_loga[3];
    int largest_active;
} sqo_timebin_params;
int64_t sqo_round_down_power_of_two(int64_t dti);
int sqo_get_timestep_bin(int64_t dti);
/* returns the number of particles print_bad_timebin would report; counts[47] */
int64_t sqo_gravity_timebins(const sqo_timebin_params * par, const double * accel, const double * gravpm, const int32_t * active,
                             const unsigned char * skip, int64_t nactive, unsigned char * timebin, int64_t * counts);
/* libgadget/timestep.cpp:386-412: largest occupied bin, push-down rule on PM steps; returns the new largest_active
 * (0 = the reference's endrun(77)); counts are updated as the reference does. */
int sqo_pushdown_bin(int64_t * counts, int largest_active, int isPM);

#ifdef __cplusplus
}
#endif
#endif
