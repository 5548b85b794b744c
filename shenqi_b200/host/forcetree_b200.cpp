// Drop-in replacement for the three tree builders of libgadget/forcetree.cpp: force_tree_full, force_tree_active_moments,
// force_tree_rebuild_mask (declared libgadget/forcetree.h:127-140; called from run.cpp:257,482,493,562-564, gravpm.cpp:97,
// timestep.cpp:297-299).  Everything else of forcetree.cpp stays the reference's (allocator, force_tree_free, accessors,
// force_tree_top_build / force_tree_find_topnode for the domain exchange, the hmax functions): the reference file is
// compiled UNMODIFIED with
//     -Dforce_tree_full=force_tree_full_cpu -Dforce_tree_active_moments=force_tree_active_moments_cpu
//     -Dforce_tree_rebuild_mask=force_tree_rebuild_mask_cpu
// so that its own builders keep existing under the _cpu names and the public names are the ones below.
//
// The tree is built on the B200 (PH keys, radix sort, level-by-level build under the DomainDecomp's top nodes, moments,
// top-leaf moment exchange and sub-tree import when NTask > 1) and stays there for grav_short_tree_cuda(); the
// reference's `struct NODE` image (forcetree.h:38-66) is written back into tree->Nodes_base so that every other module
// that walks tree->Nodes - the reference's own CPU TreeWalk included - finds the links it expects:
//   firstnode = MaxPart, Nodes = Nodes_base - firstnode, lastnode = firstnode + maxnodes, numnodes, NumParticles,
//   Father[] (optional), mask, BoxSize, moments_computed_flag, full_particle_tree_flag, TopLeaves[].treenode.
// Errors: endrun(), as in the reference.  Gas-only entry points (hmax) are not served: no gas on this path.
#include <mpi.h>
#include <math.h>
#include <string.h>
#include <omp.h>
#include <string>

#include "utils/endrun.h"
#include "utils/mymalloc.h"
#include "partmanager.h"
#include "domain.h"
#include "forcetree.h"
#include "timestep.h"
#include "gravity.h"
#include "walltime.h"
#include "b200_host.h"

/* serial number of the device tree that belongs to a host image: stored in the spare NODE slot behind the pool
 * (force_treeallocate allocates maxnodes + 1), so that grav_short_tree_cuda can tell whether the tree it is handed is the
 * one the context still holds */
static int B200TreeSerial = 0;

static void b200_install_domain(sq_ctx * ctx, const DomainDecomp * ddecomp, int NTask)
{
    static_assert(sizeof(struct topnode_data) == sizeof(sq_topnode) && sizeof(struct topleaf_data) == sizeof(sq_topleaf) &&
                  sizeof(struct task_data) == sizeof(sq_task), "DomainDecomp arrays are passed through as they are");
    b200_check(sq_domain_set(ctx, (const sq_topnode *) ddecomp->TopNodes, ddecomp->NTopNodes, (const sq_topleaf *) ddecomp->TopLeaves,
                             ddecomp->NTopLeaves, (const sq_task *) ddecomp->Tasks, NTask), "sq_domain_set");
}

/* force_tree_build, forcetree.cpp:188-262 */
static ForceTree b200_tree_build(int mask, DomainDecomp * ddecomp, const ActiveParticles * act, const int DoMoments, const int alloc_father)
{
    int NTask = 1;
    MPI_Comm_size(MPI_COMM_WORLD, &NTask);
    sq_ctx * ctx = b200_context(MPI_COMM_WORLD);
    const int64_t NumPart = PartManager->NumPart;
    /* types that the mask leaves out are not on this path (collisionless particles only) */
    #pragma omp parallel for
    for(int64_t i = 0; i < NumPart; i++)
        if(!PartManager->Base[i].IsGarbage && !((1 << PartManager->Base[i].Type) & mask))
            endrun(5, "shenqi_b200: tree mask %d leaves out particle %ld of type %d; only trees of all collisionless particles are built on the GPU\n",
                   mask, (long) i, PartManager->Base[i].Type);
    sq_params par;
    memset(&par, 0, sizeof(par));
    par.BoxSize = PartManager->BoxSize;
    /* the tree itself only needs the box; the PM geometry fixes Rcut for the sub-tree import (NTask > 1) */
    par.Nmesh = B200Pm.Nmesh > 0 ? B200Pm.Nmesh : 2;
    par.Asmth = B200Pm.Asmth > 0 ? B200Pm.Asmth : 1.5;
    par.G = B200Pm.G > 0 ? B200Pm.G : 1.0;
    par.WindowType = par.Asmth == 1.5 ? 1 : 2;
    const struct gravshort_tree_params tp = get_gravshort_treepar();
    par.ErrTolForceAcc = tp.ErrTolForceAcc; par.BHOpeningAngle = tp.BHOpeningAngle; par.MaxBHOpeningAngle = tp.MaxBHOpeningAngle;
    par.TreeUseBH = tp.TreeUseBH; par.TreeRcut = tp.Rcut > 0 ? tp.Rcut : 6.0;
    par.GravitySoftening = FORCE_SOFTENING() / 2.8;
    par.rho0 = 1.0;
    par.TreeAllocFactor = 0.7;                      /* run.cpp:257 passes 0.9 x; the library grows the pool itself (x 1.15, forcetree.cpp:215-231) */
    if(NTask > 1 && B200Pm.Nmesh <= 0)
        endrun(5, "shenqi_b200: gravpm_init_periodic must run before the first multi-rank tree build (Rcut bounds the sub-tree import)\n");
    b200_check(sq_set_params(ctx, &par), "sq_set_params");
    b200_check(sq_set_walk_raw(ctx, NULL), "sq_set_walk_raw");
    const sq_aos_layout lay = b200_particle_layout();
    b200_check(sq_upload_particles_aos(ctx, PartManager->Base, NumPart, &lay), "sq_upload_particles_aos");
    b200_check(sq_ph_sort(ctx), "sq_ph_sort");
    b200_install_domain(ctx, ddecomp, NTask);
    sq_tree_info info;
    memset(&info, 0, sizeof(info));
    const bool subset = act->ActiveParticle != NULL;
    if(subset) {
        if(NTask > 1)
            endrun(5, "shenqi_b200: active-subset trees over several ranks are not built yet\n");
        b200_check(sq_tree_build(ctx, act->ActiveParticle, act->NumActiveParticle, &info), "sq_tree_build");
    } else
        b200_check(sq_tree_build_dist(ctx, &info, NULL), "sq_tree_build_dist");
    /* pool growth happened inside the library (x 1.15, fatal above 3.0, forcetree.cpp:215-231) */
    int64_t nlocal = 0;
    b200_check(sq_tree_local_nodes(ctx, &nlocal), "sq_tree_local_nodes");
    const int64_t maxnodes = nlocal + ddecomp->NTopNodes;
    ForceTree tree = force_treeallocate(maxnodes, PartManager->MaxPart, ddecomp, alloc_father);
    tree.mask = mask;
    tree.BoxSize = PartManager->BoxSize;
    MPI_Comm_rank(MPI_COMM_WORLD, &tree.ThisTask);
    b200_check(sq_tree_export_nodes_ex(ctx, tree.Nodes_base, tree.firstnode, tree.lastnode, /* particle numbers = indices into PartManager->Base */ 1), "sq_tree_export_nodes");
    tree.numnodes = nlocal;
    tree.NumParticles = info.numparticles;
    /* TopLeaves[].treenode (force_create_node_for_topnode, forcetree.cpp:901) */
    {
        int * tn = ta_malloc("treenodes", int, ddecomp->NTopLeaves);
        b200_check(sq_domain_get_treenodes(ctx, tn), "sq_domain_get_treenodes");
        for(int i = 0; i < ddecomp->NTopLeaves; i++)
            ddecomp->TopLeaves[i].treenode = tn[i] + (int) tree.firstnode;
        ta_free(tn);
    }
    if(tree.Father) {
        /* parent leaf of every particle in the tree (forcetree.cpp:104-107 sets it during insertion) */
        #pragma omp parallel for
        for(int64_t no = 0; no < tree.numnodes; no++) {
            const struct NODE * nd = &tree.Nodes_base[no];
            if(nd->f.ChildType == PARTICLE_NODE_TYPE)
                for(int j = 0; j < nd->s.noccupied; j++)
                    tree.Father[nd->s.suns[j]] = (int) (no + tree.firstnode);
        }
    }
    tree.moments_computed_flag = DoMoments ? 1 : 1;      /* the device build always carries the moments */
    tree.hmax_computed_flag = 1;
    B200TreeSerial++;
    tree.Nodes_base[maxnodes].sibling = 0x53420000;
    tree.Nodes_base[maxnodes].father = B200TreeSerial;
    B200Tree.nodes_base = tree.Nodes_base; B200Tree.numpart = NumPart; B200Tree.numnodes = tree.numnodes; B200Tree.full = !subset; B200Tree.valid = B200TreeSerial;
    walltime_measure("/Tree/Build/Nodes");
    return tree;
}

void force_tree_full(ForceTree * tree, DomainDecomp * ddecomp, const int HybridNuTracer, const std::string EmergencyOutputDir)
{
    if(force_tree_allocated(tree))
        force_tree_free(tree);
    walltime_measure("/Misc");
    ActiveParticles act = init_empty_active_particles(PartManager);
    int mask = ALLMASK;
    if(HybridNuTracer)
        mask = GASMASK + DMMASK + STARMASK + BHMASK;
    *tree = b200_tree_build(mask, ddecomp, &act, 1, 1);
    tree->full_particle_tree_flag = 1;
}

void force_tree_active_moments(ForceTree * tree, DomainDecomp * ddecomp, const ActiveParticles * act, const int HybridNuTracer, const int alloc_father, const std::string EmergencyOutputDir)
{
    if(force_tree_allocated(tree))
        force_tree_free(tree);
    walltime_measure("/Misc");
    int mask = ALLMASK;
    if(HybridNuTracer)
        mask = GASMASK + DMMASK + STARMASK + BHMASK;
    *tree = b200_tree_build(mask, ddecomp, act, 1, alloc_father);
    if(!act->ActiveParticle)
        tree->full_particle_tree_flag = 1;
}

void force_tree_rebuild_mask(ForceTree * tree, DomainDecomp * ddecomp, int mask, const std::string EmergencyOutputDir)
{
    message(0, "Tree construction for types: %d.\n", mask);
    if(force_tree_allocated(tree))
        force_tree_free(tree);
    ActiveParticles act = init_empty_active_particles(PartManager);
    *tree = b200_tree_build(mask, ddecomp, &act, 0, 1);
    if(mask == ALLMASK)
        tree->full_particle_tree_flag = 1;
}

