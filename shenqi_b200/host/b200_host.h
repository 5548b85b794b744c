// Shared state of the B200 drop-in translation units (forcetree_b200.cpp, gravpm_b200.cpp, gravshortcuda_b200.cpp):
// one library context per process / GPU, the particle_data layout taken from the reference's own struct, and the PM
// parameters last seen (force_tree_full needs Rcut for the sub-tree import when NTask > 1).
#ifndef SHENQI_B200_HOST_H
#define SHENQI_B200_HOST_H
#include <mpi.h>
#include <stdint.h>
#include <stdlib.h>
#include <math.h>
#include <cmath>
#include <functional>
#include <memory>
#include <sstream>
#include <string>
#include "partmanager.h"
#include "forcetree.h"
#include "../../include/shenqi_b200.h"

/* the context of this rank: created on first use; NTask > 1 = one rank per GPU over NCCL (the unique id is broadcast over comm) */
sq_ctx * b200_context(MPI_Comm comm);
/* endrun() with the library's message unless rc == SQ_OK */
void b200_check(int rc, const char * what);
/* byte offsets of struct particle_data as THIS build of the reference lays it out (a DEBUG build is 176 bytes, not 160) */
sq_aos_layout b200_particle_layout(void);
/* BoxSize, Nmesh, Asmth, G as gravpm_init_periodic / gravpm_force last saw them (0 if never) */
struct b200_pm_state { double BoxSize, Asmth, G; int Nmesh; };
extern struct b200_pm_state B200Pm;

/* the device tree that force_tree_full / force_tree_active_moments of forcetree_b200.cpp left in the context, so that
 * grav_short_tree_cuda walks it instead of uploading and rebuilding (timestep.cpp:297-299 calls it once per time bin) */
struct b200_tree_cache { const struct NODE * nodes_base; int64_t numpart; int64_t numnodes; int full; int valid; /* serial, 0 = none */ };
/* 1 if `tree` is the host image of the device tree the context holds (serial number in the spare NODE slot) */
int b200_tree_is_cached(const ForceTree * tree);
extern struct b200_tree_cache B200Tree;
#endif
