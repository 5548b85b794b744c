/* TEST INFRASTRUCTURE ONLY.
 *
 * Pins SURVEY 8(f2) - gravity time-bin assignment, active sub-list, hierarchical gravity kick - to the reference's OWN
 * code: this translation unit #includes libgadget/timestep.cpp WHERE IT LIES (oracle/Makefile compiles it into
 * oracle/_ref/libshenqi_ref.so; no reference source is copied), which gives the wrappers below access to the
 * file-static functions of that file:
 *   get_timestep_gravity_dloga / grav_acceleration2   timestep.cpp:1012-1039
 *   convert_timestep_to_ti                            timestep.cpp:157-174   (TimeBinMgr::dti_from_dloga, timebinmgr.h:139-156)
 *   round_down_power_of_two                           timebinmgr.cpp:190
 *   get_timestep_bin                                  timestep.cpp:1236-1250
 *   build_active_sublist                              timestep.cpp:1375-1397
 *   do_grav_short_range_kick                          timestep.cpp:962-968
 * The loops around them (timestep.cpp:351-377, 270-287) are re-driven here call for call.  TimeBinMgr keeps its sync
 * points private; the driver sets them directly (`#define private public` around the header, test code only) so that a
 * segment crossing can be pinned without the cosmology integrals (cosmology.cpp needs boost::odeint, not in this image).
 */
#include <mpi.h>
#include <math.h>
#include <cmath>
#include <functional>
#include <memory>
#include <sstream>
#include <string>
#define private public
#include <libgadget/timebinmgr.h>
#undef private
#include <libgadget/timestep.cpp>
#include <libgadget/gravity.h>

/* ---- cosmology symbols timestep.cpp / timebinmgr.cpp reference but the pinned functions never reach ---- */
double hubble_function(const Cosmology * CP, double a) { (void) CP; (void) a; endrun(5, "ref_timestep: hubble_function is not linked\n"); return 0; }
double get_omega_nu(const _omega_nu * const omnu, const double a) { (void) omnu; (void) a; return 0; }

static TimeBinMgr RefTimeBinMgr;

extern "C" {

/* TimestepParams (timestep.cpp:28-50) and the sync points of the TimeBinMgr */
void ref_ts_setup(double ErrTolIntAccuracy, double MinSizeTimestep, const double * sync_loga, int nsync)
{
    memset(&TimestepParams, 0, sizeof(TimestepParams));
    TimestepParams.ErrTolIntAccuracy = ErrTolIntAccuracy;
    TimestepParams.MinSizeTimestep = MinSizeTimestep;
    TimestepParams.MaxSizeTimestep = 0.1;
    RefTimeBinMgr.NSyncPoints = nsync;
    RefTimeBinMgr.SyncPoints = std::unique_ptr<SyncPoint []>(new SyncPoint[nsync]);
    for(int i = 0; i < nsync; i++) {
        memset(&RefTimeBinMgr.SyncPoints[i], 0, sizeof(SyncPoint));
        RefTimeBinMgr.SyncPoints[i].loga = sync_loga[i];
    }
}

/* FORCE_SOFTENING() / 2.8 = eps (gravshort-tree2.cpp:78-93) */
void ref_ts_set_softening(double eps)
{
    struct gravshort_tree_params tp = get_gravshort_treepar();
    tp.FractionalGravitySoftening = eps;
    set_gravshort_treepar(tp);
    gravshort_set_softenings(1.0);
}

double ref_ts_loga_from_ti(int64_t ti) { return RefTimeBinMgr.loga_from_ti(ti); }

/* the bin loop of hierarchical_gravity_accelerations, timestep.cpp:351-377, over the loaded particles (GravPM, IsGarbage,
 * Swallowed from the particle table; accel = the acceleration store, N x 3).  Writes Part[].TimeBinGravity, copies it to
 * timebin[NumPart] and fills counts[TIMEBINS + 1]; returns the number of particles print_bad_timebin would report. */
int64_t ref_ts_gravity_timebins(const double * accel, const int * active, int64_t nactive, double atime, double hubble,
                                int64_t Ti_Current, int64_t dti_max, int largest_active, unsigned char * timebin, int64_t * counts)
{
    int64_t nbad = 0;
    for(int b = 0; b <= TIMEBINS; b++) counts[b] = 0;
    for(int64_t i = 0; i < nactive; i++) {
        const int pa = active ? active[i] : (int) i;
        if(Part[pa].Swallowed || Part[pa].IsGarbage)
            continue;
        const double * GravAccel = accel + 3 * (int64_t) pa;
        const double dloga_gravity = get_timestep_gravity_dloga(pa, GravAccel, atime, hubble);
        inttime_t dti_gravity = convert_timestep_to_ti(dloga_gravity, dti_max, Ti_Current, &RefTimeBinMgr);
        dti_gravity = round_down_power_of_two(dti_gravity);
        if(dti_gravity <= 1 || dti_gravity > (inttime_t) TIMEBASE)
            nbad++;
        int bin = get_timestep_bin(dti_gravity);
        if(bin > largest_active)
            bin = largest_active;
        counts[bin]++;
        Part[pa].TimeBinGravity = bin;
    }
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        timebin[i] = Part[i].TimeBinGravity;
    return nbad;
}

void ref_ts_set_timebins(const unsigned char * timebin)
{
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        Part[i].TimeBinGravity = timebin[i];
}

/* build_active_sublist, timestep.cpp:1375-1397; returns the length, list copied to out (room for nactive entries) */
int64_t ref_ts_active_sublist(const int * active, int64_t nactive, int maxtimebin, int64_t Ti_Current, int * out)
{
    ActiveParticles act = init_empty_active_particles(PartManager);
    act.ActiveParticle = (int *) active;
    act.NumActiveParticle = nactive;
    ActiveParticles sub = build_active_sublist(&act, maxtimebin, Ti_Current);
    const int64_t n = sub.NumActiveParticle;
    memcpy(out, sub.ActiveParticle, sizeof(int) * n);
    myfree(sub.ActiveParticle);
    return n;
}

/* the kick loop of apply_hierarchical_grav_kick, timestep.cpp:270-279, with the kick factor given; vel in / out N x 3 */
void ref_ts_hierarchical_kick(double * vel, const double * accel, const int * active, int64_t nactive, double gravkick)
{
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        for(int j = 0; j < 3; j++)
            Part[i].Vel[j] = vel[3 * i + j];
    for(int64_t i = 0; i < nactive; i++) {
        const int pa = active ? active[i] : (int) i;
        if(Part[pa].Swallowed || Part[pa].IsGarbage)
            continue;
        do_grav_short_range_kick(&Part[pa], accel + 3 * (int64_t) pa, gravkick);
    }
    for(int64_t i = 0; i < PartManager->NumPart; i++)
        for(int j = 0; j < 3; j++)
            vel[3 * i + j] = Part[i].Vel[j];
}

int ref_ts_get_timestep_bin(int64_t dti) { return get_timestep_bin(dti); }
int64_t ref_ts_round_down_power_of_two(int64_t dti) { return round_down_power_of_two(dti); }

} /* extern "C" */
