"""Synthetic Zel'dovich initial conditions for the benchmarks.

Lattice Ngrid^3 at (i,j,k)*Box/Ngrid (as libgenic/zeldovich.cpp:77-104), Gaussian random field with a fixed
seed (181170, benchmarks/dm-50-512/paramfile.genic:20) and a power-law-with-cutoff spectrum, displacement
psi_k = i k / k^2 delta_k (libgenic/zeldovich.cpp:292-314), normalised to a chosen rms displacement in units
of the mean spacing (z = 9: ~0.2), positions wrapped into [0, Box), equal masses Omega0 rho_crit Box^3 / N.
Pure numpy: host-side input generation, not part of the timed path."""
import numpy as np

G_INTERNAL = 43007.1          # kpc/h, 1e10 Msun/h, km/s
HUBBLE = 0.1                  # 100 km/s/Mpc in internal units
OMEGA0 = 0.2814               # benchmarks/dm-50-512/paramfile.genic


def zeldovich(ngrid, box, rms_disp=0.2, seed=181170, index=-1.0):
    """Returns pos[N,3] float64 in [0,box), mass[N] float32, ids[N] uint64 (i*Ng^2 + j*Ng + k + 1)."""
    rng = np.random.default_rng(seed)
    n = ngrid
    delta = rng.standard_normal((n, n, n))
    dk = np.fft.rfftn(delta)
    del delta
    kx = np.fft.fftfreq(n, 1.0 / n)[:, None, None]
    ky = np.fft.fftfreq(n, 1.0 / n)[None, :, None]
    kz = np.fft.rfftfreq(n, 1.0 / n)[None, None, :]
    k2 = kx * kx + ky * ky + kz * kz
    k2[0, 0, 0] = 1.0
    amp = k2 ** (index / 4.0) * np.exp(-k2 / (0.25 * n) ** 2)     # sqrt(P(k)), P ~ k^index with a cutoff at Nyquist/2
    amp[0, 0, 0] = 0.0
    dk *= amp
    disp = []
    for kk in (kx, ky, kz):
        disp.append(np.fft.irfftn(1j * kk / k2 * dk, s=(n, n, n), axes=(0, 1, 2)))
    del dk
    disp = np.stack([d.reshape(-1) for d in disp], 1)
    disp *= rms_disp * (box / n) / np.sqrt((disp ** 2).sum(1).mean() / 3.0)
    idx = np.arange(n ** 3)
    pos = np.stack([idx // (n * n), (idx // n) % n, idx % n], 1).astype(np.float64) * (box / n)
    pos += disp
    np.mod(pos, box, out=pos)
    pos[pos >= box] = 0.0
    rho_crit = 3 * HUBBLE ** 2 / (8 * np.pi * G_INTERNAL)
    mass = np.full(n ** 3, OMEGA0 * rho_crit * box ** 3 / n ** 3, np.float32)
    ids = (idx + 1).astype(np.uint64)
    return pos, mass, ids


def zeldovich_torch(ngrid, box, device, rms_disp=0.2, seed=181170, index=-1.0):
    """The same construction on a CUDA device with torch (input generation for the multi-GPU benchmark, where
    512^3 particles per rank in numpy would take minutes; torch's generator, so not the same realisation as
    zeldovich()).  Returns pos[N,3] float64 and mass[N] float32 device tensors."""
    import torch
    n = ngrid
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    dk = torch.fft.rfftn(torch.randn((n, n, n), dtype=torch.float64, device=device, generator=g))
    kx = torch.fft.fftfreq(n, 1.0 / n, dtype=torch.float64, device=device)[:, None, None]
    ky = torch.fft.fftfreq(n, 1.0 / n, dtype=torch.float64, device=device)[None, :, None]
    kz = torch.fft.rfftfreq(n, 1.0 / n, dtype=torch.float64, device=device)[None, None, :]
    k2 = kx * kx + ky * ky + kz * kz
    k2[0, 0, 0] = 1.0
    amp = k2 ** (index / 4.0) * torch.exp(-k2 / (0.25 * n) ** 2)
    amp[0, 0, 0] = 0.0
    dk *= amp
    del amp
    pos = torch.empty((n ** 3, 3), dtype=torch.float64, device=device)
    for a, kk in enumerate((kx, ky, kz)):
        pos[:, a] = torch.fft.irfftn(1j * kk / k2 * dk, s=(n, n, n)).reshape(-1)
    del dk, k2
    pos *= rms_disp * (box / n) / torch.sqrt((pos ** 2).sum(1).mean() / 3.0)
    idx = torch.arange(n ** 3, device=device)
    pos[:, 0] += (idx // (n * n)).to(torch.float64) * (box / n)
    pos[:, 1] += ((idx // n) % n).to(torch.float64) * (box / n)
    pos[:, 2] += (idx % n).to(torch.float64) * (box / n)
    del idx
    pos.remainder_(box)
    pos[pos >= box] = 0.0
    rho_crit = 3 * HUBBLE ** 2 / (8 * np.pi * G_INTERNAL)
    mass = torch.full((n ** 3,), OMEGA0 * rho_crit * box ** 3 / n ** 3, dtype=torch.float32, device=device)
    return pos, mass


def rho0():
    return OMEGA0 * 3 * HUBBLE ** 2 / (8 * np.pi * G_INTERNAL)


def clustered(ngrid, box, halo_fraction=0.5, nhalo=None, seed=181170):
    """Deterministic stand-in for evolved (z = 0-like) particle loads: the Zel'dovich
    load above with `halo_fraction` of the particles moved into Hernquist-profile halos (rho ~ 1 / (r (r + a)^3)) whose
    masses follow dn/dM ~ M^-1.9 and whose centres are drawn from the particles themselves (so halos sit in the
    overdense regions).  Deep trees, long interaction lists, strong load imbalance between groups: the stress case of
    the walk.  Returns pos, mass, ids like zeldovich()."""
    pos, mass, ids = zeldovich(ngrid, box, rms_disp=0.6, seed=seed)
    n = len(pos)
    rng = np.random.default_rng(seed + 1)
    nmove = int(halo_fraction * n)
    if nmove == 0:
        return pos, mass, ids
    nhalo = nhalo or max(8, n // 4096)
    u = rng.random(nhalo)
    mmin, mmax, al = 32.0, max(64.0, 0.02 * n), -0.9            # cumulative of M^-1.9
    hm = (mmin ** al + u * (mmax ** al - mmin ** al)) ** (1.0 / al)
    cnt = np.maximum(1, np.floor(hm * nmove / hm.sum())).astype(np.int64)
    cnt[-1] += nmove - cnt.sum()
    if cnt[-1] < 1:
        cnt[-1] = 1
    nmove = int(cnt.sum())
    moved = rng.choice(n, nmove, replace=False)
    centres = pos[rng.choice(n, nhalo, replace=False)]
    spacing = box / ngrid
    a = 0.35 * spacing * np.cbrt(cnt.astype(np.float64))        # scale radius ~ virial radius / 5 at 200 x mean density
    hid = np.repeat(np.arange(nhalo), cnt)
    q = rng.random(nmove) * 0.98                                  # enclosed-mass fraction, truncated at ~100 a
    r = a[hid] * np.sqrt(q) / (1.0 - np.sqrt(q))
    mu = 2.0 * rng.random(nmove) - 1.0
    ph = 2.0 * np.pi * rng.random(nmove)
    st = np.sqrt(1.0 - mu * mu)
    pos[moved] = centres[hid] + r[:, None] * np.stack([st * np.cos(ph), st * np.sin(ph), mu], 1)
    np.mod(pos, box, out=pos)
    pos[pos >= box] = 0.0
    return pos, mass, ids


def clustered_torch(ngrid, box, device, frac=0.5, nhalo=None, seed=181170):
    """clustered() on a CUDA device with torch (multi-GPU benchmark set-up; torch's generator, so its own realisation):
    Zel'dovich load with rms displacement 0.6 spacings, `frac` of the particles moved into Hernquist halos with
    dn/dM ~ M^-1.9 centred on particles.  Returns (pos, mass, tag); tag names generator and seed for the bench line."""
    import torch
    pos, mass = zeldovich_torch(ngrid, box, device, rms_disp=0.6, seed=seed)
    n = pos.shape[0]
    g = torch.Generator(device=device)
    g.manual_seed(seed + 1)
    nmove = int(frac * n)
    tag = "shenqi_b200.ics.clustered_torch: zeldovich rms 0.6 + %.2f of the particles in Hernquist halos dn/dM~M^-1.9, seed %d" % (frac, seed)
    if nmove == 0:
        return pos, mass, tag
    nhalo = nhalo or max(8, n // 4096)
    u = torch.rand(nhalo, dtype=torch.float64, device=device, generator=g)
    mmin, mmax, al = 32.0, max(64.0, 0.02 * n), -0.9
    hm = (mmin ** al + u * (mmax ** al - mmin ** al)) ** (1.0 / al)
    cnt = torch.clamp(torch.floor(hm * nmove / hm.sum()), min=1).to(torch.int64)
    cnt[-1] += nmove - int(cnt.sum())
    if int(cnt[-1]) < 1:
        cnt[-1] = 1
    nmove = int(cnt.sum())
    moved = torch.randperm(n, device=device, generator=g)[:nmove]
    centres = pos[torch.randint(0, n, (nhalo,), device=device, generator=g)]
    spacing = box / ngrid
    a = 0.35 * spacing * cnt.to(torch.float64) ** (1.0 / 3)
    hid = torch.repeat_interleave(torch.arange(nhalo, device=device), cnt)
    q = torch.rand(nmove, dtype=torch.float64, device=device, generator=g) * 0.98
    r = a[hid] * torch.sqrt(q) / (1.0 - torch.sqrt(q))
    mu = 2.0 * torch.rand(nmove, dtype=torch.float64, device=device, generator=g) - 1.0
    ph = 2.0 * np.pi * torch.rand(nmove, dtype=torch.float64, device=device, generator=g)
    st = torch.sqrt(1.0 - mu * mu)
    pos[moved] = centres[hid] + r[:, None] * torch.stack([st * torch.cos(ph), st * torch.sin(ph), mu], 1)
    pos.remainder_(box)
    pos[pos >= box] = 0.0
    return pos, mass, tag
