"""Synthetic SKA-like observations for the benchmark and the full-size parity tests.

No dependency on ska_sdp_datamodels / astropy: station layouts are drawn from a seeded
generator, baselines are all pairs a2 >= a1 *including*
autocorrelations (which carry zero weight, as in the reference's Visibility:
/root/reference/tests/imaging/test_imaging_ng.py:83), and uvw follow the reference's
plain-numpy convention xyz_at_latitude -> xyz_to_uvw
(/root/reference/src/ska_sdp_func_python/util/coordinate_support.py:252-272, :335-363),
restated here.
"""
import numpy as np

C_M_S = 299792458.0
SEED = 1805550721  # the reference's own test seed (tests/testing_utils.py:39)


def _disc(rng, n, radius):
    r = radius * np.sqrt(rng.uniform(0, 1, n))
    th = rng.uniform(0, 2 * np.pi, n)
    return np.stack([r * np.cos(th), r * np.sin(th), np.zeros(n)], axis=1)


def _spiral(rng, n, narms, r0, r1, jitter=0.03):
    k = np.arange(n)
    arm = k % narms
    s = (k // narms + rng.uniform(0.2, 0.8, n)) / np.ceil(n / narms)
    r = r0 * (r1 / r0) ** s
    th = 2 * np.pi * arm / narms + 1.6 * np.log(r / r0) + rng.normal(0, jitter, n)
    return np.stack([r * np.cos(th), r * np.sin(th), np.zeros(n)], axis=1)


def station_layout(name, seed=SEED):
    """East-north-up station positions [n, 3] in metres and the array latitude [rad]."""
    rng = np.random.default_rng(seed)
    if name == "LOW-full":
        enu = np.concatenate([_disc(rng, 224, 500.0), _spiral(rng, 288, 3, 600.0, 35e3)])
        lat = np.radians(-26.8)
    elif name == "LOW-core":
        enu = _disc(rng, 166, 500.0)
        lat = np.radians(-26.8)
    elif name == "MID":
        core = rng.normal(0, 300.0, (100, 3))
        core[:, 2] = 0
        enu = np.concatenate([core, _spiral(rng, 97, 3, 800.0, 75e3)])
        lat = np.radians(-30.7)
    else:
        raise ValueError(f"unknown layout {name!r}")
    return enu, lat


def enu_to_celestial_xyz(enu, lat):
    """Local (east, north, up) -> celestial xyz: x east, z towards the celestial pole."""
    e, n, u = enu[:, 0], enu[:, 1], enu[:, 2]
    colat = np.pi / 2 - lat
    y2 = -u * np.sin(colat) + n * np.cos(colat)
    z2 = u * np.cos(colat) + n * np.sin(colat)
    return np.stack([e, y2, z2], axis=1)


def xyz_to_uvw(xyz, ha, dec):
    """Rotate celestial xyz (antennas or baselines) to uvw for hour angle / declination."""
    x, y, z = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    u = x * np.cos(ha) - y * np.sin(ha)
    v0 = x * np.sin(ha) + y * np.cos(ha)
    w = z * np.sin(dec) - v0 * np.cos(dec)
    v = z * np.cos(dec) + v0 * np.sin(dec)
    return np.stack([u, v, w], axis=1)


def baseline_uvw(enu, lat, hour_angles, dec):
    """uvw [ntimes, nbaselines, 3] (metres), baselines = all pairs a2 >= a1 (autos included);
    also returns the boolean auto-correlation mask [nbaselines]."""
    xyz = enu_to_celestial_xyz(enu, lat)
    n = xyz.shape[0]
    a1, a2 = np.triu_indices(n, 0)
    out = np.empty((len(hour_angles), len(a1), 3))
    for t, ha in enumerate(hour_angles):
        ant = xyz_to_uvw(xyz, ha, dec)
        out[t] = ant[a2] - ant[a1]
    return out, a1 == a2


CONFIGS = {
    # name: layout, ntimes, HA half-range [h], nchan, f0, f1 [Hz], npix, dec [deg], eps, precision
    "C1": dict(layout="LOW-core", ntimes=2, ha_h=300.0 / 7200.0, nchan=5, f0=100e6, f1=110e6,
               npix=512, dec=-35.0, epsilon=1e-12, precision="double", cellsize=5e-4),
    "C2": dict(layout="LOW-full", ntimes=30, ha_h=1.0, nchan=64, f0=140e6, f1=160e6,
               npix=2048, dec=-35.0, epsilon=1e-5, precision="single"),
    "C3": dict(layout="MID", ntimes=120, ha_h=2.0, nchan=256, f0=950e6, f1=1050e6,
               npix=4096, dec=-45.0, epsilon=1e-5, precision="single"),
    "C4": dict(layout="LOW-full", ntimes=30, ha_h=1.0, nchan=1024, f0=120e6, f1=180e6,
               npix=8192, dec=-35.0, epsilon=1e-5, precision="single", nplanes=128),
    "C5": dict(layout="LOW-full", ntimes=30, ha_h=1.0, nchan=64, f0=140e6, f1=160e6,
               npix=16384, dec=-35.0, epsilon=1e-10, precision="double"),
}


def observation(config, seed=SEED, ntimes=None, nchan=None, row_fraction=1.0):
    """Geometry of one of the BASELINE.json configurations (optionally cut down).

    Returns a dict with uvw [nrow, 3] (metres, reference sign convention *before* the
    wrapper's u/w flip), freq [nchan], auto [nrow] (True for autocorrelation rows), npix,
    cellsize [rad], epsilon, precision, and the config name."""
    cfg = dict(CONFIGS[config]) if isinstance(config, str) else dict(config)
    name = config if isinstance(config, str) else cfg.get("name", "custom")
    enu, lat = station_layout(cfg["layout"], seed)
    nt = ntimes or cfg["ntimes"]
    nc = nchan or cfg["nchan"]
    ha = np.radians(15.0 * np.linspace(-cfg["ha_h"], cfg["ha_h"], nt))
    uvw, auto = baseline_uvw(enu, lat, ha, np.radians(cfg["dec"]))
    nbl = uvw.shape[1]
    uvw = uvw.reshape(-1, 3)
    auto = np.tile(auto, nt)
    if row_fraction < 1.0:
        rng = np.random.default_rng(seed + 7)
        keep = np.sort(rng.choice(uvw.shape[0], int(uvw.shape[0] * row_fraction), replace=False))
        uvw, auto = uvw[keep], auto[keep]
    freq = np.linspace(cfg["f0"], cfg["f1"], nc) if nc > 1 else np.array([cfg["f0"]])
    uvmax = np.abs(uvw[:, :2]).max() * cfg["f1"] / C_M_S
    cellsize = cfg.get("cellsize") or 1.0 / (4.0 * uvmax)
    if cfg.get("nplanes"):
        # w rescaled by a constant so that the reference model (sigma = 1.5,
        # W = 7, dw = 1/(2 sigma max|n-1|)) needs ceil((wmax-wmin)/dw) + W = nplanes w-planes
        xm = cellsize * cfg["npix"] / 2.0
        r2 = 2.0 * xm * xm
        nm1 = r2 / (np.sqrt(1.0 - r2) + 1.0)
        dw = 0.5 / (1.5 * nm1)
        wmax = np.abs(uvw[:, 2]).max() * cfg["f1"] / C_M_S
        uvw = uvw.copy()
        uvw[:, 2] *= (cfg["nplanes"] - 7 - 0.5) * dw / wmax
    return dict(name=name, uvw=np.ascontiguousarray(uvw), freq=freq, auto=auto, nbaselines=nbl,
                ntimes=nt, npix=cfg["npix"], cellsize=float(cellsize), epsilon=cfg["epsilon"],
                precision=cfg["precision"], nplanes=cfg.get("nplanes"))


def flip_for_gridder(uvw):
    """The wrapper's u and w sign flip (ng.py:80-85, :209-213)."""
    out = np.array(uvw, dtype=np.float64, copy=True)
    out[:, 0] *= -1.0
    out[:, 2] *= -1.0
    return out


def point_source_vis(uvw, freq, npix, cellsize, nsrc=32, seed=SEED + 1, do_wstacking=True):
    """Exact visibilities (gridder/ducc0 convention, i.e. for *flipped* uvw) of nsrc random
    point sources on pixel centres; returns (ms [nrow, nchan] c128, model image [npix, npix])."""
    rng = np.random.default_rng(seed)
    ix = rng.integers(npix // 8, npix - npix // 8, nsrc)
    iy = rng.integers(npix // 8, npix - npix // 8, nsrc)
    flux = rng.uniform(0.5, 2.0, nsrc)
    model = np.zeros((npix, npix))
    np.add.at(model, (ix, iy), flux)
    x = (ix - npix // 2) * cellsize
    y = (iy - npix // 2) * cellsize
    r2 = x * x + y * y
    nm1 = -r2 / (np.sqrt(1.0 - r2) + 1.0) if do_wstacking else np.zeros_like(r2)
    n = nm1 + 1.0
    ms = np.zeros((uvw.shape[0], len(freq)), dtype=np.complex128)
    fl = np.asarray(freq) / C_M_S
    for k in range(nsrc):
        ph = uvw[:, 0] * x[k] + uvw[:, 1] * y[k] - uvw[:, 2] * nm1[k]
        ms += (flux[k] / n[k]) * np.exp(-2j * np.pi * np.outer(ph, fl))
    return ms, model
