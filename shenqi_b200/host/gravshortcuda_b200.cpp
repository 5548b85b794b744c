// Drop-in replacement for the reference's libgadget/gravshortcuda.cu.
//
// The reference's short-range gravity entry point, grav_short_tree() (libgadget/gravshort-tree2.cpp:141-172),
// hands the walk to
//     void grav_short_tree_cuda(const ActiveParticles * act, ForceTree * tree, GravTreeParams * priv,
//                               GravTreeOutput * output, particle_data * const parts,
//                               const size_t MaxExportBufferBytes, MPI_Comm comm)
// when the code is built with USE_CUDA and run with UseGPU (declared libgadget/gravshort2.hpp:440-443, defined
// in libgadget/gravshortcuda.cu:20-27 as a TreeWalkGPU instantiation).  This translation unit defines the
// same symbol on top of the C ABI of libshenqi_b200.so, so run.cpp / timestep.cpp / the tests call the B200
// path without any change: compile it INSTEAD of gravshortcuda.cu.
//
// Contract kept:
//   reads   act->ActiveParticle (NULL = all [0, NumActiveParticle)), parts[].Pos/Mass/FullTreeGravAccel/GravPM,
//           every field of *priv (GravTreeParams, gravshort2.hpp:21-55) incl. the float split tables,
//           tree->full_particle_tree_flag / tree->BoxSize
//   writes  output->Accel[i][0..2] (already x G) for every active i; on a full-particle tree also
//           parts[i].FullTreeGravAccel and parts[i].Potential (= tree potential with the self terms of
//           GravTreeOutput::postprocess, gravshort2.hpp:88-107; the primary reduce ASSIGNS the potential)
//   errors  endrun() (abort), like every reference module; there is no CPU fallback.
// With forcetree_b200.cpp linked, the device tree built by force_tree_full / force_tree_active_moments for this host
// tree is walked as it is (also for NTask > 1: domain top nodes, exchanged moments, imported sub-trees); otherwise the
// device tree is rebuilt from the particles (single rank), tree->Nodes of the host tree is not read.
#include <mpi.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "gravshort2.hpp"
#include "utils/endrun.h"
#include "b200_host.h"

static sq_walk_stats B200LastStats;

extern "C" const sq_walk_stats * shenqi_b200_last_walk_stats(void) { return &B200LastStats; }

void
grav_short_tree_cuda(const ActiveParticles * act, ForceTree * tree, GravTreeParams * priv, GravTreeOutput * output, particle_data * const parts, const size_t MaxExportBufferBytes, MPI_Comm comm)
{
    int NTask = 1;
    MPI_Comm_size(comm, &NTask);
    sq_ctx * B200Ctx = b200_context(comm);
    /* The device tree that forcetree_b200.cpp's force_tree_full / force_tree_active_moments built for THIS host tree is
     * walked as it is (top nodes of the domain, moments exchanged, remote sub-trees imported when NTask > 1): no second
     * upload, sort or build.  A host tree from the reference's own forcetree.cpp is rebuilt on the device from the particles. */
    const bool cached = b200_tree_is_cached(tree) != 0;
    if(NTask > 1 && !cached)
        endrun(5, "shenqi_b200: with NTask = %d the tree must come from forcetree_b200.cpp (it carries the domain's top nodes and the imported sub-trees)\n", NTask);

    const int64_t NumPart = PartManager->NumPart;
    /* GravShortTable keeps its tables private; apply_short_range_window() at the knots returns them exactly
     * (gravity.h:50-60: the interpolation weights are 1 and 0 there). */
    const double table_dx = 0.029354207436397859;   /* gravshort-tree2.cpp:48 */
    std::vector<float> ftab(NGRAVTAB), ptab(NGRAVTAB);
    for(int i = 0; i < NGRAVTAB; i++) {
        double fac = 1, pot = 1;
        const double r = (i + 1e-9) * table_dx * priv->cellsize;
        if(0 != priv->gravtab.apply_short_range_window(r, &fac, &pot, priv->cellsize)) {
            /* last knot: not reachable directly (tabindex >= NGRAVTAB-1 means "outside"); it is the other end of
             * the last interpolation interval: value at 510.5 = (t[510] + t[511]) / 2 */
            fac = pot = 1;
            priv->gravtab.apply_short_range_window((i - 0.5) * table_dx * priv->cellsize, &fac, &pot, priv->cellsize);
            fac = 2 * fac - ftab[i - 1];
            pot = 2 * pot - ptab[i - 1];
        }
        ftab[i] = (float) fac;
        ptab[i] = (float) pot;
    }
    sq_params par;
    memset(&par, 0, sizeof(par));
    par.BoxSize = priv->BoxSize;
    par.Nmesh = (int) lround(priv->BoxSize / priv->cellsize);
    par.Asmth = 1.5; par.WindowType = 1;
    par.G = priv->G;
    par.ErrTolForceAcc = priv->ErrTolForceAcc;
    par.TreeUseBH = priv->TreeUseBH;
    par.GravitySoftening = priv->ForceSoftening / 2.8;
    par.rho0 = priv->cbrtrho0 * priv->cbrtrho0 * priv->cbrtrho0;
    par.TreeAllocFactor = 0.9;
    b200_check(sq_set_params(B200Ctx, &par), "sq_set_params");
    sq_walk_raw raw;
    raw.BoxSize = priv->BoxSize; raw.cellsize = priv->cellsize; raw.Rcut = priv->Rcut; raw.G = priv->G;
    raw.cbrtrho0 = priv->cbrtrho0; raw.ForceSoftening = priv->ForceSoftening; raw.ErrTolForceAcc = priv->ErrTolForceAcc;
    raw.BHOpeningAngle2 = priv->BHOpeningAngle2; raw.TreeUseBH = priv->TreeUseBH;
    raw.ntab = NGRAVTAB; raw.force_table = ftab.data(); raw.pot_table = ptab.data(); raw.table_dx = table_dx;
    b200_check(sq_set_walk_raw(B200Ctx, &raw), "sq_set_walk_raw");

    const int * active = act->ActiveParticle;
    const int64_t nactive = act->NumActiveParticle;
    const bool all = (active == NULL && nactive == NumPart);
    sq_tree_info info;
    memset(&info, 0, sizeof(info));
    if(cached) {
        /* only |a_old| = |FullTreeGravAccel + GravPM| / G may have changed since the build (gravshort2.hpp:111-121): 8 bytes per particle */
        std::vector<double> oldacc((size_t) NumPart + 1);
        #pragma omp parallel for
        for(int64_t i = 0; i < NumPart; i++) {
            double s = 0;
            for(int d = 0; d < 3; d++) { const double ax = parts[i].FullTreeGravAccel[d] + parts[i].GravPM[d]; s += ax * ax; }
            oldacc[i] = sqrt(s) / priv->G;
        }
        b200_check(sq_set_oldacc(B200Ctx, oldacc.data()), "sq_set_oldacc");
        info.numnodes = tree->numnodes;
    } else {
        /* particle_data is read in place (layout from the reference's own struct); |a_old| is formed on the device */
        const sq_aos_layout lay = b200_particle_layout();
        b200_check(sq_upload_particles_aos(B200Ctx, parts, NumPart, &lay), "sq_upload_particles_aos");
        b200_check(sq_ph_sort(B200Ctx), "sq_ph_sort");
        b200_check(sq_domain_set(B200Ctx, NULL, 0, NULL, 0, NULL, 0), "sq_domain_set");
        /* the tree holds what the host tree holds: every particle on a full-particle tree (PM steps), else the active set */
        if(tree->full_particle_tree_flag)
            b200_check(sq_tree_build(B200Ctx, NULL, 0, &info), "sq_tree_build");
        else {
            std::vector<int32_t> idx;
            const int32_t * lst = active;
            if(!lst) { idx.resize(nactive); for(int64_t i = 0; i < nactive; i++) idx[i] = (int32_t) i; lst = idx.data(); }
            b200_check(sq_tree_build(B200Ctx, lst, nactive, &info), "sq_tree_build");
        }
    }
    std::vector<double> accel((size_t) NumPart * 3), pot;
    /* inactive entries must keep the caller's values */
    static_assert(sizeof(MyFloat) == sizeof(double), "LOW_PRECISION=double expected (Makefile.rules:6)");
    memcpy(accel.data(), output->Accel, sizeof(double) * 3 * (size_t) NumPart);
    if(output->update_potential) pot.assign((size_t) NumPart, 0.0);
    std::vector<int32_t> idx;
    const int32_t * lst = NULL;
    if(!all) {
        lst = active;
        if(!lst) { idx.resize(nactive); for(int64_t i = 0; i < nactive; i++) idx[i] = (int32_t) i; lst = idx.data(); }
    }
    b200_check(sq_tree_walk(B200Ctx, lst, all ? 0 : nactive, accel.data(), output->update_potential ? pot.data() : NULL, &B200LastStats), "sq_tree_walk");
    /* GravTreeResult::reduce<PRIMARY> + GravTreeOutput::postprocess (gravshort2.hpp:136-146, 88-107) */
    #pragma omp parallel for
    for(int64_t k = 0; k < nactive; k++) {
        const int64_t i = active ? active[k] : k;
        for(int d = 0; d < 3; d++)
            output->Accel[i][d] = accel[3 * i + d];
        if(output->update_potential) {
            for(int d = 0; d < 3; d++)
                parts[i].FullTreeGravAccel[d] = accel[3 * i + d];
            parts[i].Potential = pot[i];
        }
    }
    message(0, "shenqi_b200 GRAVTREE: %ld targets, %ld interactions, walk kernel %.3f ms, %ld nodes\n",
            (long) B200LastStats.ntargets, (long) B200LastStats.interactions, B200LastStats.kernel_ms, (long) info.numnodes);
}
