// Context, error plumbing and particle layout shared by the B200 drop-in translation units (b200_host.h).
#include <mpi.h>
#include <stddef.h>
#include <string.h>
#include "utils/endrun.h"
#include "b200_host.h"

struct b200_pm_state B200Pm = {0, 0, 0, 0};
struct b200_tree_cache B200Tree = {NULL, 0, 0, 0, 0};
static sq_ctx * B200Ctx = NULL;

void b200_check(int rc, const char * what)
{
    if(rc != SQ_OK)
        endrun(5, "shenqi_b200: %s failed (%d): %s\n", what, rc, sq_last_error());
}

sq_ctx * b200_context(MPI_Comm comm)
{
    if(B200Ctx)
        return B200Ctx;
    int NTask = 1, ThisTask = 0;
    MPI_Comm_size(comm, &NTask);
    MPI_Comm_rank(comm, &ThisTask);
    if(NTask == 1) {
        b200_check(sq_ctx_create(&B200Ctx, 0), "sq_ctx_create");
        return B200Ctx;
    }
    /* one rank per GPU (README.rst:67): rank 0 makes the NCCL id, everybody gets it over the host communicator */
    unsigned char id[SQ_NCCL_UNIQUE_ID_BYTES];
    memset(id, 0, sizeof(id));
    if(ThisTask == 0)
        b200_check(sq_comm_unique_id(id), "sq_comm_unique_id");
    MPI_Bcast(id, sizeof(id), MPI_BYTE, 0, comm);
    /* ranks of one node take consecutive devices; sq_ctx_create_dist refuses a device that does not exist */
    int ndev = 1;
    const char * vis = getenv("SHENQI_B200_GPUS_PER_NODE");
    if(vis) ndev = atoi(vis) > 0 ? atoi(vis) : 1;
    else ndev = NTask < 8 ? NTask : 8;
    b200_check(sq_ctx_create_dist(&B200Ctx, ThisTask % ndev, id, ThisTask, NTask), "sq_ctx_create_dist");
    return B200Ctx;
}

sq_aos_layout b200_particle_layout(void)
{
    sq_aos_layout lay;
    memset(&lay, 0, sizeof(lay));
    lay.stride = sizeof(struct particle_data);
    lay.off_pos = offsetof(struct particle_data, Pos);
    lay.off_mass = offsetof(struct particle_data, Mass);
    /* the anonymous flag struct starts right behind PI: bit 0 IsGarbage, bit 1 Swallowed, byte + 3 Type (partmanager.h:18-31) */
    lay.off_flags = offsetof(struct particle_data, PI) + sizeof(int);
    lay.off_fulltreegravaccel = offsetof(struct particle_data, FullTreeGravAccel);
    lay.off_gravpm = offsetof(struct particle_data, GravPM);
    lay.off_id = offsetof(struct particle_data, ID);
    lay.off_potential = offsetof(struct particle_data, Potential);
    /* the Type byte must be where the kernel reads it */
    struct particle_data probe;
    memset(&probe, 0, sizeof(probe));
    probe.Type = 5; probe.Swallowed = 1;
    const unsigned char * b = (const unsigned char *) &probe + lay.off_flags;
    if(b[3] != 5 || !(b[0] & 2))
        endrun(5, "shenqi_b200: unexpected layout of the particle_data flag word\n");
    return lay;
}

int b200_tree_is_cached(const ForceTree * tree)
{
    if(!B200Tree.valid || tree->Nodes_base != B200Tree.nodes_base || tree->numnodes != B200Tree.numnodes || PartManager->NumPart != B200Tree.numpart)
        return 0;
    const struct NODE * spare = &tree->Nodes_base[tree->lastnode - tree->firstnode];
    return spare->sibling == 0x53420000 && spare->father == B200Tree.valid;
}
