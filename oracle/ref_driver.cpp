/* TEST INFRASTRUCTURE ONLY.
 *
 * C entry points around the UNMODIFIED reference sources (compiled where they lie under
 * /root/reference by oracle/Makefile into oracle/_ref/libshenqi_ref.so).  This file contains no
 * reference code: it only fills the reference's own structs (PartManager, DomainDecomp, PetaPM,
 * gravshort_tree_params) and calls the reference's own functions
 *   PEANO / peano_hilbert_key      libgadget/utils/peano.h:15, peano.cpp:144
 *   slots_gc_sorted                libgadget/slotsmanager.cpp:419
 *   force_tree_full                libgadget/forcetree.cpp:110
 *   grav_short_tree                libgadget/gravshort-tree2.cpp:141
 *   GravLocalTreeWalk::visit       libgadget/gravshort2.hpp:227   (interaction counts)
 *   grav_short_pair                libgadget/gravshort-pair.c:22
 *   drift_all_particles            libgadget/drift.cpp:81
 * the way libgadget/tests/test_forcetree.cpp:294-331 and test_gravity.cpp:197-247 do
 * (single rank, trivial domain).  Only tests/, smoke() and bench.py's CPU-baseline leg load it.
 */
#include <mpi.h>
#include <string.h>
#include <math.h>
#include <omp.h>
#include <string>

#include <libgadget/utils/mymalloc.h>
#include <libgadget/utils/endrun.h>
#include <libgadget/utils/peano.h>
#include <libgadget/partmanager.h>
#include <libgadget/slotsmanager.h>
#include <libgadget/walltime.h>
#include <libgadget/domain.h>
#include <libgadget/forcetree.h>
#include <libgadget/gravity.h>
#include <libgadget/gravity_old.h>
#include <libgadget/petapm.h>
#include <libgadget/timestep.h>
#include <libgadget/drift.h>
#include <libgadget/gravshort2.hpp>
#include <libgadget/cosmology.h>

/* ---- symbols the path references but which live in reference files we do not build ---- */
/* (init_empty_active_particles and is_timebin_active now come from the reference's own timestep.cpp, compiled through oracle/ref_timestep.cpp) */

/* libgadget/bhdynfric.h:22, consulted by drift.cpp:32 for Type 5 only; black holes are not built here */
int BHGetRepositionEnabled(void) { return 0; }

static struct ClockTable RefClocks;
static DomainDecomp RefDomain;
static ForceTree RefTree;
static int ref_initialised = 0;

static void ref_trivial_domain(DomainDecomp * ddecomp)
{
    memset(ddecomp, 0, sizeof(*ddecomp));
    ddecomp->domain_allocated_flag = 1;
    ddecomp->NTopNodes = 1;
    ddecomp->NTopLeaves = 1;
    ddecomp->TopNodes = (struct topnode_data *) calloc(1, sizeof(struct topnode_data));
#ifdef REF_USE_MANAGED
    /* TopLeaves are dereferenced by the reference's toptree kernels */
    ddecomp->TopLeaves = mymanagedmalloc("TopLeaves", struct topleaf_data, 1);
    memset(ddecomp->TopLeaves, 0, sizeof(struct topleaf_data));
#else
    ddecomp->TopLeaves = (struct topleaf_data *) calloc(1, sizeof(struct topleaf_data));
#endif
    ddecomp->TopNodes[0].Daughter = -1;
    ddecomp->TopNodes[0].Leaf = 0;
    ddecomp->TopNodes[0].StartKey = 0;
    ddecomp->TopNodes[0].Shift = BITS_PER_DIMENSION * 3;
    ddecomp->TopLeaves[0].Task = 0;
    ddecomp->Tasks = (struct task_data *) calloc(1, sizeof(struct task_data));
    ddecomp->Tasks[0].StartLeaf = 0;
    ddecomp->Tasks[0].EndLeaf = 1;
}

extern "C" {

int ref_init(void)
{
    if(ref_initialised)
        return 0;
    int thr;
    MPI_Init_thread(NULL, NULL, MPI_THREAD_FUNNELED, &thr);
    init_endrun(0);
#ifdef REF_USE_MANAGED
    /* the reference's own GPU walk (tests/dropin libshenqi_refgpu.so) reads particles, nodes and parameters from managed memory */
    allocator_malloc_init(A_MAIN, "MAIN", 1);
    allocator_malloc_init(A_TEMP, "TEMP", 1);
#else
    allocator_malloc_init(A_MAIN, "MAIN", 0);
    allocator_malloc_init(A_TEMP, "TEMP", 0);
#endif
    walltime_init(&RefClocks);
    memset(PartManager, 0, sizeof(PartManager[0]));
    memset(SlotsManager, 0, sizeof(SlotsManager[0]));
    memset(&RefTree, 0, sizeof(RefTree));
    ref_trivial_domain(&RefDomain);
    ref_initialised = 1;
    return 0;
}

int ref_sizeof_particle(void) { return (int) sizeof(struct particle_data); }
int ref_sizeof_node(void) { return (int) sizeof(struct NODE); }

uint64_t ref_peano_hilbert_key(int x, int y, int z, int bits) { return peano_hilbert_key(x, y, z, bits); }

void ref_peano_keys(const double * pos, int64_t n, double box, uint64_t * keys)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < n; i++)
        keys[i] = PEANO(pos + 3 * i, box);
}

/* Load n particles (type 1 unless types given) into the reference's particle manager. */
void ref_set_particles(const double * pos, const float * mass, const uint64_t * ids, const unsigned char * types, int64_t n, double box)
{
    ref_init();
    if(force_tree_allocated(&RefTree))
        force_tree_free(&RefTree);
    if(PartManager->Base)
        myfree(PartManager->Base);
    memset(PartManager, 0, sizeof(PartManager[0]));
    particle_alloc_memory(PartManager, box, n);
    memset(PartManager->Base, 0, sizeof(struct particle_data) * n);
    #pragma omp parallel for
    for(int64_t i = 0; i < n; i++) {
        struct particle_data * p = &PartManager->Base[i];
        for(int j = 0; j < 3; j++)
            p->Pos[j] = pos[3 * i + j];
        p->Mass = mass[i];
        p->ID = ids ? ids[i] : (uint64_t) (i + 1);
        p->Type = types ? types[i] : 1;
    }
    PartManager->NumPart = n;
    PartManager->BoxSize = box;
}

int64_t ref_num_particles(void) { return PartManager->NumPart; }

void ref_set_accels(const double * fulltree, const double * gravpm)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        for(int j = 0; j < 3; j++) {
            if(fulltree) PartManager->Base[i].FullTreeGravAccel[j] = fulltree[3 * i + j];
            if(gravpm) PartManager->Base[i].GravPM[j] = gravpm[3 * i + j];
        }
}

void ref_get_particles(double * pos, float * mass, uint64_t * ids, double * fulltree, double * potential)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++) {
        const struct particle_data * p = &PartManager->Base[i];
        for(int j = 0; j < 3; j++) {
            if(pos) pos[3 * i + j] = p->Pos[j];
            if(fulltree) fulltree[3 * i + j] = p->FullTreeGravAccel[j];
        }
        if(mass) mass[i] = p->Mass;
        if(ids) ids[i] = p->ID;
        if(potential) potential[i] = p->Potential;
    }
}

/* drift_all_particles (drift.cpp:81-99) on the loaded particles with the given velocities; flags: bit0 IsGarbage, bit1 Swallowed */
void ref_drift_all(const double * vel, const unsigned char * flags, double ddrift, const double * random_shift)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++) {
        struct particle_data * p = &PartManager->Base[i];
        for(int j = 0; j < 3; j++)
            p->Vel[j] = vel[3 * i + j];
        p->IsGarbage = flags ? (flags[i] & 1) : 0;
        p->Swallowed = flags ? ((flags[i] >> 1) & 1) : 0;
        p->Ti_drift = 0;
    }
    drift_all_particles(0, 1, ddrift, random_shift);
}

/* raw access to the reference's AoS, for boundary-layout tests */
void * ref_particle_base(void) { return PartManager->Base; }

/* slots_gc_sorted: PH key + sort by (type,key) + permute */
double ref_sort_particles(void)
{
    double t0 = MPI_Wtime();
    slots_gc_sorted(PartManager, SlotsManager);
    return MPI_Wtime() - t0;
}

/* A multi-leaf DomainDecomp on the single rank (every top leaf owned by task 0): force_tree_full then builds the forced
 * top-level nodes of force_create_node_for_topnode (forcetree.cpp:868-933) above the sub-trees.  Arrays in the layout of
 * struct topnode_data (24 B: StartKey, Daughter, Shift, Leaf) and topleaf_data (12 B); ntop == 0 restores the trivial domain. */
void ref_set_domain(const void * topnodes, int ntop, const void * topleaves, int nleaves)
{
    ref_init();
    if(force_tree_allocated(&RefTree))
        force_tree_free(&RefTree);
    free(RefDomain.TopNodes); free(RefDomain.TopLeaves); free(RefDomain.Tasks);
    if(ntop <= 0) { ref_trivial_domain(&RefDomain); return; }
    memset(&RefDomain, 0, sizeof(RefDomain));
    RefDomain.domain_allocated_flag = 1;
    RefDomain.NTopNodes = ntop;
    RefDomain.NTopLeaves = nleaves;
    RefDomain.TopNodes = (struct topnode_data *) calloc(ntop, sizeof(struct topnode_data));
    RefDomain.TopLeaves = (struct topleaf_data *) calloc(nleaves, sizeof(struct topleaf_data));
    RefDomain.Tasks = (struct task_data *) calloc(1, sizeof(struct task_data));
    memcpy(RefDomain.TopNodes, topnodes, ntop * sizeof(struct topnode_data));
    memcpy(RefDomain.TopLeaves, topleaves, nleaves * sizeof(struct topleaf_data));
    for(int i = 0; i < nleaves; i++)
        RefDomain.TopLeaves[i].Task = 0;
    RefDomain.Tasks[0].StartLeaf = 0;
    RefDomain.Tasks[0].EndLeaf = nleaves;
}
int ref_sizeof_topnode(void) { return (int) sizeof(struct topnode_data); }
int ref_sizeof_topleaf(void) { return (int) sizeof(struct topleaf_data); }
/* TopLeaves[].treenode after the build, relative to firstnode */
void ref_domain_treenodes(int * out)
{
    for(int i = 0; i < RefDomain.NTopLeaves; i++)
        out[i] = RefDomain.TopLeaves[i].treenode - (int) RefTree.firstnode;
}

/* force_tree_full with the trivial single-rank domain. Returns numnodes; seconds in *secs. */
int64_t ref_tree_build(double treeallocfactor, double * secs)
{
    init_forcetree_params(treeallocfactor);
    /* Part[i].TopLeaf as the domain decomposition leaves it (domain.cpp: domain_get_topleaf of the particle's key) */
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        PartManager->Base[i].TopLeaf = domain_get_topleaf(PEANO(PartManager->Base[i].Pos, PartManager->BoxSize), RefDomain.TopNodes);
    double t0 = MPI_Wtime();
    force_tree_full(&RefTree, &RefDomain, 0, "");
    if(secs) *secs = MPI_Wtime() - t0;
    return RefTree.numnodes;
}

void ref_tree_free(void)
{
    if(force_tree_allocated(&RefTree))
        force_tree_free(&RefTree);
}

int64_t ref_tree_firstnode(void) { return RefTree.firstnode; }
int64_t ref_tree_numnodes(void) { return RefTree.numnodes; }
int64_t ref_tree_numparticles(void) { return RefTree.NumParticles; }
/* copy the NODE image [firstnode, firstnode+numnodes) out as raw bytes */
void ref_tree_copy_nodes(void * out)
{
    memcpy(out, RefTree.Nodes_base, sizeof(struct NODE) * RefTree.numnodes);
}

static PetaPM ref_make_pm(int Nmesh, double Asmth, double G)
{
    PetaPM pm;
    memset(&pm, 0, sizeof(pm));
    pm.Nmesh = Nmesh;
    pm.Asmth = Asmth;
    pm.G = G;
    pm.BoxSize = PartManager->BoxSize;
    pm.CellSize = PartManager->BoxSize / Nmesh;
    return pm;
}

static struct gravshort_tree_params ref_make_treepar(double ErrTolForceAcc, double BHOpeningAngle, double MaxBHOpeningAngle,
        int TreeUseBH, double Rcut, double FracSoft, int windowtype)
{
    struct gravshort_tree_params tp;
    memset(&tp, 0, sizeof(tp));
    tp.ErrTolForceAcc = ErrTolForceAcc;
    tp.BHOpeningAngle = BHOpeningAngle;
    tp.MaxBHOpeningAngle = MaxBHOpeningAngle;
    tp.TreeUseBH = TreeUseBH;
    tp.Rcut = Rcut;
    tp.FractionalGravitySoftening = FracSoft;
    tp.MaxExportBufferBytes = 1024 * 1024 * 1024L;
    tp.ShortRangeForceWindowType = (enum ShortRangeForceWindowType) windowtype;
    return tp;
}

/* grav_short_tree on all particles (tree must be built). accel: N x 3 (already x G). Returns seconds. */
/* UseGPU argument of grav_short_tree: honoured only by builds with USE_CUDA (tests/dropin links the B200 hook) */
static int RefUseGPU = 0;
void ref_set_use_gpu(int v) { RefUseGPU = v; }

double ref_grav_short_tree(int Nmesh, double Asmth, double G, double ErrTolForceAcc, double BHOpeningAngle,
        double MaxBHOpeningAngle, int TreeUseBH, double Rcut, double FracSoft, double MeanSeparation,
        int windowtype, double rho0, double * accel)
{
    PetaPM pm = ref_make_pm(Nmesh, Asmth, G);
    set_gravshort_treepar(ref_make_treepar(ErrTolForceAcc, BHOpeningAngle, MaxBHOpeningAngle, TreeUseBH, Rcut, FracSoft, windowtype));
    gravshort_set_softenings(MeanSeparation);
    ActiveParticles act = init_empty_active_particles(PartManager);
    double t0 = MPI_Wtime();
    grav_short_tree(&act, &pm, &RefTree, (MyFloat (*)[3]) accel, rho0, 0, RefUseGPU != 0);
    return MPI_Wtime() - t0;
}

/* Per-particle interaction counts of the reference's own local walk (gravshort2.hpp:227-322),
 * with the same parameters; does not modify any particle. Returns the total. */
int64_t ref_grav_count_interactions(int Nmesh, double Asmth, double G, double ErrTolForceAcc, double BHOpeningAngle,
        double MaxBHOpeningAngle, int TreeUseBH, double Rcut, double FracSoft, double MeanSeparation,
        int windowtype, int64_t * counts)
{
    PetaPM pm = ref_make_pm(Nmesh, Asmth, G);
    struct gravshort_tree_params tp = ref_make_treepar(ErrTolForceAcc, BHOpeningAngle, MaxBHOpeningAngle, TreeUseBH, Rcut, FracSoft, windowtype);
    set_gravshort_treepar(tp);
    gravshort_set_softenings(MeanSeparation);
    GravShortTable gravtab(tp.ShortRangeForceWindowType, Asmth);
    GravTreeParams priv(tp, 0, 1.0, &pm, RefTree.BoxSize, gravtab);
    int64_t total = 0;
    #pragma omp parallel for schedule(dynamic, 64) reduction(+: total)
    for(int64_t i = 0; i < PartManager->NumPart; i++) {
        GravTreeQuery input(PartManager->Base[i], NULL, RefTree.firstnode, priv);
        GravTreeResult output(input);
        GravLocalTreeWalk lv(RefTree.Nodes, input);
        const int64_t n = lv.visit<TREEWALK_PRIMARY>(input, &output, priv, PartManager->Base);
        if(counts) counts[i] = n;
        total += n;
    }
    return total;
}

/* grav_short_tree over an explicit active list (particle indices of the loaded, sorted set), as the hierarchical
 * stepping calls it (timestep.cpp:297-299).  accel: NumPart x 3, only the rows of active particles are written. */
double ref_grav_short_tree_active(int Nmesh, double Asmth, double G, double ErrTolForceAcc, double BHOpeningAngle,
        double MaxBHOpeningAngle, int TreeUseBH, double Rcut, double FracSoft, double MeanSeparation,
        int windowtype, double rho0, const int * active, int64_t nactive, double * accel)
{
    PetaPM pm = ref_make_pm(Nmesh, Asmth, G);
    set_gravshort_treepar(ref_make_treepar(ErrTolForceAcc, BHOpeningAngle, MaxBHOpeningAngle, TreeUseBH, Rcut, FracSoft, windowtype));
    gravshort_set_softenings(MeanSeparation);
    ActiveParticles act = init_empty_active_particles(PartManager);
    act.ActiveParticle = (int *) active;
    act.NumActiveParticle = nactive;
    double t0 = MPI_Wtime();
    grav_short_tree(&act, &pm, &RefTree, (MyFloat (*)[3]) accel, rho0, 0, RefUseGPU != 0);
    return MPI_Wtime() - t0;
}

/* interaction counts of the reference's local walk for the listed particles only; counts[k] belongs to active[k] */
int64_t ref_grav_count_interactions_active(int Nmesh, double Asmth, double G, double ErrTolForceAcc, double BHOpeningAngle,
        double MaxBHOpeningAngle, int TreeUseBH, double Rcut, double FracSoft, double MeanSeparation,
        int windowtype, const int * active, int64_t nactive, int64_t * counts)
{
    PetaPM pm = ref_make_pm(Nmesh, Asmth, G);
    struct gravshort_tree_params tp = ref_make_treepar(ErrTolForceAcc, BHOpeningAngle, MaxBHOpeningAngle, TreeUseBH, Rcut, FracSoft, windowtype);
    set_gravshort_treepar(tp);
    gravshort_set_softenings(MeanSeparation);
    GravShortTable gravtab(tp.ShortRangeForceWindowType, Asmth);
    GravTreeParams priv(tp, 0, 1.0, &pm, RefTree.BoxSize, gravtab);
    int64_t total = 0;
    #pragma omp parallel for schedule(dynamic, 64) reduction(+: total)
    for(int64_t k = 0; k < nactive; k++) {
        const int i = active[k];
        GravTreeQuery input(PartManager->Base[i], NULL, RefTree.firstnode, priv);
        GravTreeResult output(input);
        GravLocalTreeWalk lv(RefTree.Nodes, input);
        const int64_t n = lv.visit<TREEWALK_PRIMARY>(input, &output, priv, PartManager->Base);
        if(counts) counts[k] = n;
        total += n;
    }
    return total;
}

/* keys of the loaded particles in their current (after ref_sort_particles: sorted) order */
void ref_current_keys(uint64_t * keys)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        keys[i] = PEANO(PartManager->Base[i].Pos, PartManager->BoxSize);
}

/* grav_short_pair: direct sum over r <= Rcut; result lands in FullTreeGravAccel/Potential (tree is full). */
double ref_grav_short_pair(int Nmesh, double Asmth, double G, double Rcut, double FracSoft, double MeanSeparation,
        int windowtype, double rho0, double * accel)
{
    PetaPM pm = ref_make_pm(Nmesh, Asmth, G);
    struct gravshort_tree_params tp = ref_make_treepar(0.002, 0.175, 0.9, 0, Rcut, FracSoft, windowtype);
    set_gravshort_treepar(tp);
    set_gravshort_treepar_old(tp);
    gravshort_set_softenings(MeanSeparation);
    ActiveParticles act = init_empty_active_particles(PartManager);
    double t0 = MPI_Wtime();
    grav_short_pair(&act, &pm, &RefTree, Rcut, rho0);
    double dt = MPI_Wtime() - t0;
    if(accel) {
        #pragma omp parallel for
        for(int64_t i = 0; i < PartManager->NumPart; i++)
            for(int j = 0; j < 3; j++)
                accel[3 * i + j] = PartManager->Base[i].FullTreeGravAccel[j];
    }
    return dt;
}

#ifdef REF_HAVE_GRAVPM
/* gravpm_init_periodic + gravpm_force (gravity.h:71, 86) on the loaded particles.  Two builds link a gravpm
 * implementation: tests/dropin (shenqi_b200/host/gravpm_b200.cpp behind the reference's symbols) and, with
 * REF_REFERENCE_PM, oracle/_ref itself: the reference's OWN gravpm.cpp + petapm.cpp + powerspectrum.cpp, unmodified,
 * with only the FFT library behind heFFTe replaced (ref_shim/heffte.h).  The power spectrum file
 * (powerspectrum_save, powerspectrum.cpp:90-118) goes to the directory set with ref_set_power_dir. */
static std::string RefPowerDir(".");
#ifdef REF_REFERENCE_PM
static void ref_pm_module_init(void);
#endif
void ref_set_power_dir(const char * dir) { RefPowerDir = dir ? dir : "."; }
void ref_gravpm_force(int Nmesh, double Asmth, double G)
{
    static PetaPM pm;
#ifdef REF_REFERENCE_PM
    ref_pm_module_init();
    /* what run.cpp:257 and the domain decomposition leave behind before gravpm_force builds its tree (gravpm.cpp:97) */
    init_forcetree_params(0.9);
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        PartManager->Base[i].TopLeaf = domain_get_topleaf(PEANO(PartManager->Base[i].Pos, PartManager->BoxSize), RefDomain.TopNodes);
#endif
    gravpm_init_periodic(&pm, PartManager->BoxSize, Asmth, Nmesh, G);
    Cosmology CP;
    memset(&CP, 0, sizeof(CP));
    CP.Omega0 = 0.3; CP.OmegaLambda = 0.7; CP.HubbleParam = 0.7;
    gravpm_force(&pm, &RefDomain, &CP, 0.1, 3.085678e21, RefPowerDir, 0.01);
#ifdef REF_REFERENCE_PM
    petapm_destroy(&pm);
#endif
}
#endif
#ifdef REF_REFERENCE_PM
static void ref_pm_module_init(void)
{
    static int done = 0;
    if(!done) { petapm_module_init(omp_get_max_threads(), 0); done = 1; }
}
/* petapm_init + petapm_fft_r2c + petapm_fft_c2r (petapm.h:141-146, petapm.cpp:49-71, 197-332) on a caller-supplied mesh:
 * real_in[x][y][z] (z fastest: the real-space region's strides) -> cplx_out in the fourier region's own order, indexed
 * (y, z, x) with x fastest (petapm.cpp:265-272), Nmesh^2 (Nmesh/2+1) complex numbers; real_back = c2r(cplx_out), unscaled.
 * Also returns the two regions' sizes and strides (6 + 6 entries) as petapm_init left them. */
void ref_petapm_fft(int Nmesh, const double * real_in, double * cplx_out, double * real_back, int64_t * regions)
{
    ref_init();
    ref_pm_module_init();
    static PetaPM pm;
    petapm_init(&pm, 1.0, 1.25, Nmesh, 1.0, MPI_COMM_WORLD);
    const int64_t nreal = (int64_t) Nmesh * Nmesh * Nmesh, ncplx = (int64_t) Nmesh * Nmesh * (Nmesh / 2 + 1);
    double * real = mymalloc("fftreal", double, pm.priv->fftsize);
    petapm_complex * cplx = (petapm_complex *) mymalloc("fftcplx", double, pm.priv->fftsize);
    memcpy(real, real_in, sizeof(double) * nreal);
    petapm_fft_r2c(&pm, real, cplx);
    memcpy(cplx_out, cplx, sizeof(double) * 2 * ncplx);
    petapm_fft_c2r(&pm, cplx, real);
    memcpy(real_back, real, sizeof(double) * nreal);
    PetaPMRegion * rr = petapm_get_real_region(&pm), * fr = petapm_get_fourier_region(&pm);
    for(int k = 0; k < 3; k++) {
        regions[k] = rr->size[k]; regions[3 + k] = rr->strides[k];
        regions[6 + k] = fr->size[k]; regions[9 + k] = fr->strides[k];
    }
    myfree(cplx);
    myfree(real);
    petapm_destroy(&pm);
}
/* powerspectrum_add_mode (gravpm.cpp:325-356) + powerspectrum_sum (powerspectrum.cpp:52-88) on caller-supplied modes:
 * kpos[m][3] signed mesh wave numbers, value[m][2].  Returns the number of non-empty bins; kk / power / nmodes hold
 * Nmesh entries each. */
int ref_powerspectrum_bin(int Nmesh, double BoxSize_in_cm, int64_t nmodes_in, const int * kpos, const double * value, const double * invwindow,
                          double * kk, double * power, int64_t * nmodes, double * norm)
{
    Power ps;
    memset((void *) &ps, 0, sizeof(ps));
    powerspectrum_alloc(&ps, Nmesh, 1, 0, BoxSize_in_cm);
    for(int64_t m = 0; m < nmodes_in; m++) {
        const int * kp = kpos + 3 * m;
        const int64_t k2 = (int64_t) kp[0] * kp[0] + (int64_t) kp[1] * kp[1] + (int64_t) kp[2] * kp[2];
        petapm_complex v = {value[2 * m], value[2 * m + 1]};
        powerspectrum_add_mode(&ps, k2, kp, &v, invwindow ? invwindow[m] : 1.0, Nmesh);
    }
    powerspectrum_sum(&ps);
    for(int i = 0; i < ps.nonzero; i++) { kk[i] = ps.kk[i]; power[i] = ps.Power[i]; nmodes[i] = ps.Nmodes[i]; }
    *norm = ps.Norm;
    const int nz = ps.nonzero;
    powerspectrum_free(&ps);
    return nz;
}
#endif
void ref_set_potential(const double * pot)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        PartManager->Base[i].Potential = pot[i];
}
void ref_get_gravpm(double * gravpm)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        for(int j = 0; j < 3; j++)
            gravpm[3 * i + j] = PartManager->Base[i].GravPM[j];
}
void ref_set_flags(const unsigned char * flags, const unsigned char * types)
{
    #pragma omp parallel for
    for(int64_t i = 0; i < PartManager->NumPart; i++) {
        PartManager->Base[i].IsGarbage = flags[i] & 1;
        PartManager->Base[i].Swallowed = (flags[i] >> 1) & 1;
        if(types) PartManager->Base[i].Type = types[i];
    }
}
int64_t ref_tree_numparticles2(void) { return RefTree.NumParticles; }
int ref_tree_flags(void) { return RefTree.moments_computed_flag | (RefTree.full_particle_tree_flag << 1) | ((RefTree.Father != NULL) << 2); }
/* Father[i] - firstnode for every loaded particle (needs a tree built with alloc_father) */
void ref_tree_fathers(int * out)
{
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        out[i] = RefTree.Father ? RefTree.Father[i] - (int) RefTree.firstnode : -1;
}

double ref_force_softening(void) { return FORCE_SOFTENING(); }
int ref_omp_threads(void) { return omp_get_max_threads(); }
/* torchrun exports OMP_NUM_THREADS=1: the benchmark's CPU arm asks for every host core explicitly */
void ref_set_threads(int n) { if(n > 0) omp_set_num_threads(n); }

} /* extern "C" */
