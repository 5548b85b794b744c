"""The reference's OWN particle-mesh code as the checker of the PM restatement.

Round 1 pinned the PM only through the force-error bounds of libgadget/tests/test_gravity.cpp, because petapm.cpp needs
FFTW3 + heFFTe, which this image does not have.  oracle/_ref now compiles the reference's gravpm.cpp, petapm.cpp and
powerspectrum.cpp UNMODIFIED against two stand-in headers (oracle/ref_shim/heffte.h, fftw3.h: one rank, a plain
double-precision FFT of our own behind heFFTe's API), so paint (petapm.cpp:1132-1183, 1304-1310), regions
(gravpm.cpp:121-281), pencil layout (petapm.cpp:692-1107), transfer functions (gravpm.cpp:295-487) and read-out
(gravpm.cpp:489-500) run as the reference wrote them.  These tests compare
  * the stand-in FFT, through the reference's petapm_fft_r2c / petapm_fft_c2r, with numpy's;
  * gravpm_force of the reference with the restatement (o_pm_force: oracle/shenqi_oracle.c + numpy FFT) that the GPU
    parity tests use as their checker: accelerations and potentials to 1e-12 of the largest value;
  * the same with a many-leaf DomainDecomp (many regions, the pencil exchange between overlapping regions);
  * swallowed particles (gravpm.cpp:171-178): neither painted nor read out;
  * the whole do_force_test of test_gravity.cpp (:197-247) by reference code alone against the direct sum, with the
    reference's bounds, and against the restatement's result for the same flow;
  * P(k) binning (gravpm.cpp:325-356, powerspectrum.cpp:52-88) with a numpy restatement.
No GPU, no product code: this file checks the checker."""
import os
import numpy as np
import pytest

from helpers import (ROOT, ref, ptr, c_i64, c_dbl, c_int, quiet_stdout, random_clustered, default_params, o_keys, o_pm_force, o_cic_paint,
                     o_direct27, r_load, r_tree, r_walk, reference_gravity_test_samples, reference_gravity_test_metric, has_reference_pm,
                     r_gravpm_force, UNIT_CM)


pytestmark = pytest.mark.skipif(not has_reference_pm(), reason="oracle/_ref was built without the reference's PM")


def close(a, b, tol):
    return np.abs(a - b).max() <= tol * np.abs(b).max()


# ------------------------------------------------------------------------------------------ FFT stand-in
@pytest.mark.parametrize("nmesh", [8, 12, 14, 30, 48])
def test_fft_standin_through_petapm_fft(nmesh):
    """petapm_fft_r2c / _c2r (petapm.cpp:49-71): unscaled, real region z fastest, fourier region (y, z, x) with x
    fastest (petapm.cpp:265-272); sizes with factors 2, 3, 5, 7."""
    R = ref()
    N = nmesh
    a = np.random.default_rng(N).standard_normal((N, N, N))
    c = np.zeros((N, N // 2 + 1, N), np.complex128)
    b = np.zeros((N, N, N)); reg = np.zeros(12, np.int64)
    with quiet_stdout():
        R.ref_petapm_fft(c_int(N), ptr(a), ptr(c), ptr(b), ptr(reg))
    assert list(reg[:6]) == [N, N, N, N * N, N, 1]
    assert list(reg[6:]) == [N, N // 2 + 1, N, (N // 2 + 1) * N, N, 1]
    f = np.fft.rfftn(a)                                           # [x][y][kz]
    assert close(c.transpose(2, 0, 1), f, 1e-14)
    assert close(b, a * float(N) ** 3, 1e-14)


def test_fft_standin_c2r_ignores_imaginary_parts_of_self_conjugate_modes():
    """c2r of a half-mesh that is not exactly Hermitian on the kz = 0 and Nyquist planes (what the force transfer leaves
    behind) is the FFTW / cuFFT / numpy.irfftn answer."""
    R = ref()
    N = 12
    rng = np.random.default_rng(1)
    a = rng.standard_normal((N, N, N))
    c = np.zeros((N, N // 2 + 1, N), np.complex128); b = np.zeros((N, N, N)); reg = np.zeros(12, np.int64)
    with quiet_stdout():
        R.ref_petapm_fft(c_int(N), ptr(a), ptr(c), ptr(b), ptr(reg))
    # numpy's own answer for the same spectrum
    back = np.fft.irfftn(c.transpose(2, 0, 1), s=(N, N, N), axes=(0, 1, 2)) * float(N) ** 3
    assert close(b, back, 1e-14)


# ------------------------------------------------------------------------------------------ gravpm_force
CASES = [("uniform", 20000, 32, 1.25), ("clustered", 30000, 48, 1.5), ("clustered", 12000, 40, 1.5), ("lattice", 4096, 24, 1.25)]


def sample(kind, n, box, seed=5):
    if kind == "uniform":
        return np.random.default_rng(seed).random((n, 3)) * box
    if kind == "lattice":
        return reference_gravity_test_samples()[0][1] * (box / 8.0)
    return random_clustered(n, box, seed)


@pytest.mark.parametrize("kind,n,nmesh,asmth", CASES)
def test_reference_gravpm_force_equals_restatement(kind, n, nmesh, asmth):
    box, G = 8.0, 43007.1
    pos = sample(kind, n, box)
    mass = (1.0 + 0.25 * (np.arange(len(pos)) % 3)).astype(np.float32)
    pot0 = np.random.default_rng(2).standard_normal(len(pos))       # gravpm_force ADDS to Potential (gravpm.cpp:489-491)
    pot, g, _ = r_gravpm_force(pos, mass, box, nmesh, asmth, G, pot0=pot0)
    opot, oacc = o_pm_force(pos, mass, box, nmesh, asmth, G)
    assert close(g, oacc, 1e-12), np.abs(g - oacc).max() / np.abs(oacc).max()
    assert close(pot - pot0, opot, 1e-12)


def test_reference_pm_regions_of_a_many_leaf_domain():
    """every top leaf (or node <= 24 cells) becomes a region with its own mesh buffer, regions overlap by the stencil
    and are summed through the pencil layout (gravpm.cpp:121-200, petapm.cpp:692-1061): the forces do not depend on the
    decomposition, which is why the product paints into one slab without regions."""
    import shenqi_b200 as sq
    box, G, nmesh, n = 8.0, 43007.1, 48, 25000
    pos = random_clustered(n, box, 9)
    mass = np.ones(n, np.float32)
    keys = o_keys(pos, box)
    cnt = np.bincount((keys >> np.uint64(63 - 9)).astype(np.int64), minlength=512)
    dom = sq.domain_from_histogram(cnt, 1, 16)
    assert dom.ntopleaves >= 16
    pot, g, _ = r_gravpm_force(pos, mass, box, nmesh, 1.5, G, dom=dom)
    opot, oacc = o_pm_force(pos, mass, box, nmesh, 1.5, G)
    assert close(g, oacc, 1e-12) and close(pot, opot, 1e-12)


def test_reference_pm_skips_swallowed_particles():
    """Swallowed black holes (Type 5: the only particles the tree leaves out for it, forcetree.cpp:804) get
    RegionInd = -2 (gravpm.cpp:171-178): they are neither painted nor read out, GravPM stays 0 - the behaviour the
    product's skip byte reproduces (sq_pm.cu paints and reads out the n_valid prefix only)."""
    box, G, nmesh, n = 8.0, 43007.1, 32, 8000
    pos = random_clustered(n, box, 4)
    mass = np.ones(n, np.float32)
    flags = np.zeros(n, np.uint8)
    flags[np.random.default_rng(0).choice(n, 500, replace=False)] = 2
    pot, g, _ = r_gravpm_force(pos, mass, box, nmesh, 1.25, G, flags=flags, types=np.where(flags == 2, 5, 1))
    live = flags == 0
    opot, oacc = o_pm_force(np.ascontiguousarray(pos[live]), np.ascontiguousarray(mass[live]), box, nmesh, 1.25, G)
    assert np.all(g[~live] == 0) and np.all(pot[~live] == 0)
    assert close(g[live], oacc, 1e-12) and close(pot[live], opot, 1e-12)


# ------------------------------------------------------------------------------------------ test_gravity.cpp by reference code
@pytest.mark.parametrize("case", [1, 2])
def test_reference_code_alone_meets_its_gravity_test_bounds(case):
    """do_force_test (test_gravity.cpp:197-247) with the reference's own gravpm_force, force_tree_full and
    grav_short_tree (Barnes-Hut call, then the relative criterion with MaxBHOpeningAngle left 0) against the 27-image
    direct sum: maxerr < 3 ErrTol, meanerr < 0.8 ErrTol (:165-166).  The stand-in FFT is the only non-reference part."""
    from test_oracle_cpu import oracle_reference_gravity_test
    name, pos, _ = reference_gravity_test_samples()[case]
    box, nmesh, n = 8.0, 48, len(pos)
    par = default_params(box, nmesh, n, TreeUseBH=1, TreeRcut=7.0, GravitySoftening=(box / 16.0) / 30.0)
    mass = np.ones(n, np.float32)
    _, gpm, _ = r_gravpm_force(pos, mass, box, nmesh, 1.5, par["G"])        # leaves GravPM in the particle table
    with quiet_stdout():
        r_tree(0.9)
        r_walk(par, box / 16.0)
        a1, _ = r_walk(dict(par, TreeUseBH=0, MaxBHOpeningAngle=0.0), box / 16.0)
        ref().ref_tree_free()
    tot = a1 + gpm
    d = o_direct27(pos, None, box, par["G"], 2.8 * par["GravitySoftening"])
    maxerr, meanerr = reference_gravity_test_metric(tot, d)
    assert maxerr < 3 * 0.002 and meanerr < 0.8 * 0.002, (name, maxerr, meanerr)
    otot, _ = oracle_reference_gravity_test(pos)
    assert close(tot, otot, 1e-9), np.abs(tot - otot).max() / np.abs(otot).max()


# ------------------------------------------------------------------------------------------ P(k) (row f4)
def o_powerspectrum(value, kpos, nmesh, box_cm, invwindow=None):
    """numpy restatement of powerspectrum_add_mode (gravpm.cpp:325-356) + powerspectrum_sum (powerspectrum.cpp:52-88):
    log-spaced bins floor(binsperunit * log(k2) / 2), weight 1 on the kz = 0 and Nyquist planes, else 2; the zero mode is
    the normalisation; output in Mpc/h units, empty bins removed.  Returns (k, P, Nmodes, Norm)."""
    kpos = np.asarray(kpos, np.int64)
    k2 = (kpos ** 2).sum(1)
    m = (np.asarray(value) * np.conj(value)).real
    norm = m[k2 == 0][-1] if np.any(k2 == 0) else 0.0
    sel = k2 > 0
    binsperunit = (nmesh - 1) / np.log(np.sqrt(3.) * nmesh / 2.0)
    kint = np.floor(binsperunit * np.log(k2[sel].astype(np.float64)) / 2.).astype(np.int64)
    ok = kint < nmesh
    w = np.where((kpos[sel, 2] == 0) | (kpos[sel, 2] == nmesh // 2), 1, 2)[ok]
    iw = np.ones(sel.sum()) if invwindow is None else np.asarray(invwindow)[sel]
    kint = kint[ok]
    P = np.bincount(kint, weights=w * m[sel][ok] * iw[ok] ** 2, minlength=nmesh)
    K = np.bincount(kint, weights=w * np.sqrt(k2[sel][ok].astype(np.float64)), minlength=nmesh)
    Nm = np.bincount(kint, weights=w, minlength=nmesh).astype(np.int64)
    nz = Nm > 0
    box_mpc = box_cm / 3.085678e24
    return K[nz] / Nm[nz] * 2 * np.pi / box_mpc, P[nz] / Nm[nz] / norm * box_mpc ** 3, Nm[nz], norm


def half_mesh_modes(nmesh):
    f = np.fft.fftfreq(nmesh, 1.0 / nmesh).astype(np.int64)
    kx, ky, kz = np.meshgrid(f, f, np.arange(nmesh // 2 + 1), indexing="ij")
    return np.stack([kx.ravel(), ky.ravel(), kz.ravel()], 1)


def test_powerspectrum_binning_restatement_equals_reference():
    R = ref()
    nmesh, box_cm = 24, 8.0 * UNIT_CM
    kpos = half_mesh_modes(nmesh).astype(np.int32)
    rng = np.random.default_rng(6)
    val = rng.standard_normal(len(kpos)) + 1j * rng.standard_normal(len(kpos))
    iw = 1.0 + rng.random(len(kpos))
    kk = np.zeros(nmesh); pw = np.zeros(nmesh); nm = np.zeros(nmesh, np.int64); norm = c_dbl()
    import ctypes
    nz = R.ref_powerspectrum_bin(c_int(nmesh), c_dbl(box_cm), c_i64(len(kpos)), ptr(np.ascontiguousarray(kpos)), ptr(np.ascontiguousarray(val)), ptr(iw),
                                 ptr(kk), ptr(pw), ptr(nm), ctypes.byref(norm))
    ok, op, on, onorm = o_powerspectrum(val, kpos, nmesh, box_cm, iw)
    assert nz == len(ok) and np.array_equal(nm[:nz], on)
    assert int(on.sum()) == nmesh ** 3 - 1                           # every mode but the mean, counted with its multiplicity
    assert abs(norm.value - onorm) <= 1e-15 * onorm
    assert np.allclose(kk[:nz], ok, rtol=1e-13) and np.allclose(pw[:nz], op, rtol=1e-12)


def test_power_spectrum_file_of_gravpm_force():
    """the file powerspectrum_save writes at the end of gravpm_force (gravpm.cpp:108-112): the deconvolved |delta_k|^2 of
    the painted mesh in the reference's bins (printed with %g)."""
    box, G, nmesh, n = 8.0, 43007.1, 32, 20000
    pos = random_clustered(n, box, 12)
    mass = np.ones(n, np.float32)
    _, _, rows = r_gravpm_force(pos, mass, box, nmesh, 1.25, G)
    rho_k = np.fft.rfftn(o_cic_paint(pos, mass, box, nmesh))
    kpos = half_mesh_modes(nmesh)
    s = np.sinc(kpos / float(nmesh))                                  # numpy sinc(x) = sin(pi x) / (pi x)
    iw = np.prod(1.0 / s ** 2, axis=1)                                # gravpm.cpp:391-395: the CIC deconvolution, once per axis
    ok, op, on, _ = o_powerspectrum(rho_k.ravel(), kpos, nmesh, box * UNIT_CM, iw)
    assert len(rows) == len(ok) and np.array_equal(rows[:, 2].astype(np.int64), on)
    assert np.allclose(rows[:, 0], ok, rtol=2e-5) and np.allclose(rows[:, 1], op, rtol=2e-5)
