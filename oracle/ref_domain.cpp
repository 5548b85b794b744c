/* TEST INFRASTRUCTURE ONLY.
 *
 * Pins the host side of SURVEY 8(f1) - the balanced assignment of top leaves to tasks by measured cost, and the shape
 * of a DomainDecomp the reference itself produces - to the reference's OWN code: this translation unit #includes
 * libgadget/domain.cpp WHERE IT LIES (oracle/Makefile compiles it into oracle/_ref/libshenqi_ref.so; no reference source
 * is copied), which gives the wrappers below access to its file-static functions:
 *   domain_assign_topleaves_balanced   domain.cpp:620-778   (purely local: reads only the task COUNT)
 *   domain_set_task_leafs              domain.cpp:781-809
 *   domain_decompose_full              domain.cpp:174-277   (one rank: the top tree it refines from the loaded particles)
 * The task count of the purely local functions comes from MPI_Comm_size, which the one-rank mpi.h stand-in lets this
 * translation unit override (shim_mpi_fake_size); nothing that communicates runs at a faked size.
 */
#include <mpi.h>
#include <math.h>
#include <cmath>
#include <functional>
#include <memory>
#include <sstream>
#include <string>
#include <libgadget/domain.cpp>

/* libgadget/bhdynfric.h: tree mask of the black-hole dynamic friction, consulted by domain_maintain only */
int blackhole_dynfric_treemask(void) { return 0; }

extern "C" {

/* domain_assign_topleaves_balanced + domain_set_task_leafs on a caller-supplied top tree.
 * topnodes: ntopnodes records (24 B, domain.h:12-18); leaf_topnode[k]: top node of leaf k (any order); cost[k]: its cost.
 * Out, per leaf AFTER the reference's re-sort by (Task, Key): task_out[k], topnode_out[k]; tasks_out[2 * t] = StartLeaf,
 * [2 * t + 1] = EndLeaf.  Returns 0. */
int ref_domain_assign_balanced(int ntask, int nsegment_per_task, int ntopnodes, const struct topnode_data * topnodes, int ntopleaves,
                               const int * leaf_topnode, const int64_t * cost, int * task_out, int * topnode_out, int * tasks_out)
{
    DomainDecomp dd;
    memset(&dd, 0, sizeof(dd));
    dd.DomainComm = MPI_COMM_WORLD;
    dd.NTopNodes = ntopnodes;
    dd.NTopLeaves = ntopleaves;
    dd.TopNodes = (struct topnode_data *) malloc(sizeof(struct topnode_data) * ntopnodes);
    memcpy(dd.TopNodes, topnodes, sizeof(struct topnode_data) * ntopnodes);
    dd.TopLeaves = (struct topleaf_data *) calloc(ntopleaves + 1, sizeof(struct topleaf_data));
    for(int k = 0; k < ntopleaves; k++)
        dd.TopLeaves[k].topnode = leaf_topnode[k];
    int64_t * c = (int64_t *) malloc(sizeof(int64_t) * ntopleaves);
    memcpy(c, cost, sizeof(int64_t) * ntopleaves);
    shim_mpi_fake_size = ntask;
    domain_assign_topleaves_balanced(&dd, c, nsegment_per_task);
    struct task_data * tasks = domain_set_task_leafs(&dd);
    shim_mpi_fake_size = 0;
    for(int k = 0; k < ntopleaves; k++) {
        task_out[k] = dd.TopLeaves[k].Task;
        topnode_out[k] = dd.TopLeaves[k].topnode;
    }
    for(int t = 0; t < ntask; t++) {
        tasks_out[2 * t] = tasks[t].StartLeaf;
        tasks_out[2 * t + 1] = tasks[t].EndLeaf;
    }
    myfree(tasks);
    free(c); free(dd.TopLeaves); free(dd.TopNodes);
    return 0;
}

/* domain_decompose_full (domain.cpp:174-277) on the loaded particles, one rank: refines the top tree until no leaf holds
 * more than 1 / (DomainOverDecompositionFactor * NTask) of the load, assigns the leaves, exchanges (nothing to do on one
 * rank) and sorts the particles.  Copies out at most max_topnodes / max_topleaves records; returns NTopNodes, writes
 * NTopLeaves. */
int ref_domain_decompose_full(int overdecomposition, double topnode_alloc_factor, struct topnode_data * topnodes, int max_topnodes,
                              struct topleaf_data * topleaves, int max_topleaves, int * ntopleaves)
{
    DomainParams dp;
    memset(&dp, 0, sizeof(dp));
    dp.DomainOverDecompositionFactor = overdecomposition;
    dp.TopNodeAllocFactor = topnode_alloc_factor;
    dp.SetAsideFactor = 1.;
    set_domain_par(dp);
    static DomainDecomp dd;
    if(dd.domain_allocated_flag)
        domain_free(&dd);
    memset(&dd, 0, sizeof(dd));
    domain_decompose_full(&dd, MPI_COMM_WORLD);
    const int nn = dd.NTopNodes, nl = dd.NTopLeaves;
    memcpy(topnodes, dd.TopNodes, sizeof(struct topnode_data) * (nn < max_topnodes ? nn : max_topnodes));
    memcpy(topleaves, dd.TopLeaves, sizeof(struct topleaf_data) * (nl < max_topleaves ? nl : max_topleaves));
    *ntopleaves = nl;
    return nn;
}

/* Part[i].TopLeaf of the loaded particles, as domain_decompose_full stored it (domain.cpp:237-246) */
void ref_domain_get_topleaf(int * topleaf)
{
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        topleaf[i] = PartManager->Base[i].TopLeaf;
}

}   /* extern "C" */
