// Drop-in replacement for libgadget/gravpm.cpp's entry points gravpm_init_periodic and gravpm_force (declared
// libgadget/gravity.h:71, 86; called from run.cpp:347, 538, 836, 877 and the tests): compile this file INSTEAD of
// gravpm.cpp + petapm.cpp for collisionless runs.
//
// Contract kept (gravpm.cpp:60-120, 489-500):
//   zeroes Part[i].GravPM, then fills it with the long-range PM acceleration;
//   ADDS the PM potential to Part[i].Potential (readout_potential, gravpm.cpp:489-491);
//   every particle takes part (the hybrid-neutrino `active` filter and the massive-neutrino k-space hooks are not on
//   this path: MassiveNuLinRespOn / hybrid tracers are refused);
//   pm->ps: the matter power spectrum that potential_transfer bins (gravpm.cpp:430) is SURVEY 8(f4) and is not produced
//   here: pm->ps is left untouched and no powerspectrum file is written.
// The reference builds a force tree only to cut the particles into mesh regions (gravpm.cpp:97, 121-281); one GPU paints
// its whole mesh, so no tree is built here.  Errors: endrun().
#include <mpi.h>
#include <math.h>
#include <string.h>
#include <string>

#include "utils/endrun.h"
#include "utils/mymalloc.h"
#include "partmanager.h"
#include "domain.h"
#include "cosmology.h"
#include "petapm.h"
#include "gravity.h"
#include "walltime.h"
#include "b200_host.h"

void gravpm_init_periodic(PetaPM * pm, double BoxSize, double Asmth, int Nmesh, double G)
{
    /* the fields of petapm_init (petapm.cpp:197-215) that callers read; the FFT plans live in the library context */
    memset(pm, 0, sizeof(*pm));
    pm->comm = MPI_COMM_WORLD;
    pm->Nmesh = Nmesh;
    pm->Asmth = Asmth;
    pm->BoxSize = BoxSize;
    pm->CellSize = BoxSize / Nmesh;
    pm->G = G;
    B200Pm.BoxSize = BoxSize; B200Pm.Asmth = Asmth; B200Pm.G = G; B200Pm.Nmesh = Nmesh;
}

void gravpm_force(PetaPM * pm, DomainDecomp * ddecomp, Cosmology * CP, double Time, double UnitLength_in_cm, const std::string& PowerOutputDir, double TimeIC)
{
    int NTask = 1;
    MPI_Comm_size(MPI_COMM_WORLD, &NTask);
    if(NTask > 1)
        endrun(5, "shenqi_b200: the multi-rank PM (x-slabs over NCCL) is driven by shenqi_b200.dist, not by this translation unit yet; NTask = %d\n", NTask);
    if(CP && CP->MassiveNuLinRespOn)
        endrun(5, "shenqi_b200: MassiveNuLinRespOn needs the k-space neutrino hooks of gravpm.cpp:76-80, which are not on the GPU path\n");
    sq_ctx * ctx = b200_context(MPI_COMM_WORLD);
    B200Pm.BoxSize = pm->BoxSize; B200Pm.Asmth = pm->Asmth; B200Pm.G = pm->G; B200Pm.Nmesh = pm->Nmesh;
    const int64_t NumPart = PartManager->NumPart;
    sq_params par;
    memset(&par, 0, sizeof(par));
    par.BoxSize = pm->BoxSize; par.Nmesh = pm->Nmesh; par.Asmth = pm->Asmth; par.G = pm->G;
    const struct gravshort_tree_params tp = get_gravshort_treepar();
    par.ErrTolForceAcc = tp.ErrTolForceAcc; par.BHOpeningAngle = tp.BHOpeningAngle; par.MaxBHOpeningAngle = tp.MaxBHOpeningAngle;
    par.TreeUseBH = tp.TreeUseBH; par.TreeRcut = tp.Rcut > 0 ? tp.Rcut : 6.0;
    par.GravitySoftening = FORCE_SOFTENING() / 2.8;
    par.WindowType = pm->Asmth == 1.5 ? 1 : 2;
    par.rho0 = 1.0; par.TreeAllocFactor = 0.7;
    b200_check(sq_set_params(ctx, &par), "sq_set_params");
    const sq_aos_layout lay = b200_particle_layout();
    b200_check(sq_upload_particles_aos(ctx, PartManager->Base, NumPart, &lay), "sq_upload_particles_aos");
    b200_check(sq_ph_sort(ctx), "sq_ph_sort");
    B200Tree.valid = 0;                 /* the context's particle set was replaced: a device tree from before is gone */
    double * gpm = mymalloc("b200gravpm", double, 3 * (NumPart + 1));
    double * pot = mymalloc("b200pmpot", double, NumPart + 1);
    b200_check(sq_pm_force(ctx, gpm, pot), "sq_pm_force");
    #pragma omp parallel for
    for(int64_t i = 0; i < NumPart; i++) {
        struct particle_data * p = &PartManager->Base[i];
        /* gravpm.cpp:85-89 zeroes GravPM first; readout_force_x/y/z add, readout_potential adds to Potential (:489-500) */
        for(int k = 0; k < 3; k++)
            p->GravPM[k] = gpm[3 * i + k];
        p->Potential += pot[i];
    }
    myfree(pot);
    myfree(gpm);
    walltime_measure("/PMgrav/Misc");
}
