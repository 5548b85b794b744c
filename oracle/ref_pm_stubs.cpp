/* TEST INFRASTRUCTURE ONLY.
 *
 * The reference's gravpm.cpp references five functions of its cosmology / massive-neutrino modules.  Those modules are
 * outside the hot path (out of scope) and need Boost (odeint, tanh_sinh, barycentric_rational), which this
 * image does not have, so oracle/_ref links these stand-ins instead.  With Cosmology::MassiveNuLinRespOn == 0 and no
 * hybrid neutrinos - the only setting the oracle runs - gravpm_force (gravpm.cpp:59-118) calls just GrowthFactor, and
 * only to print the D1 column of the power spectrum file (powerspectrum.cpp:111-115).
 */
#include <stdio.h>
#include <stdlib.h>
#include <libgadget/cosmology.h>
#include <libgadget/neutrinos_lra.h>

/* cosmology.h: linear growth factor between two scale factors; matter-dominated stand-in D ~ a */
double GrowthFactor(Cosmology * CP, double astart, double aend)
{
    (void) CP;
    return astart / aend;
}

/* neutrinos_lra.h: are neutrino particles passive tracers at this time?  never in the oracle's runs */
int hybrid_nu_tracer(const Cosmology * CP, double atime)
{
    (void) CP; (void) atime;
    return 0;
}

void delta_nu_from_power(Power * PowerSpectrum, Cosmology * CP, const double Time, const double TimeIC)
{
    (void) PowerSpectrum; (void) CP; (void) Time; (void) TimeIC;
    fprintf(stderr, "oracle: delta_nu_from_power reached - the linear-response neutrino module is not part of the oracle build\n");
    abort();
}

double get_omega_nu_nopart(const _omega_nu * const omnu, const double a)
{
    (void) omnu; (void) a;
    return 0.;
}

/* neutrinos_lra.h:69: only reached when a neutrino spline was allocated (gravpm.cpp:114-115) */
void powerspectrum_nu_save(struct _powerspectrum * PowerSpectrum, const std::string OutputDir, const std::string filename, const double Time)
{
    (void) PowerSpectrum; (void) OutputDir; (void) filename; (void) Time;
}
