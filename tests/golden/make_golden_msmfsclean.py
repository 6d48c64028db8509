#!/usr/bin/env python
"""Golden vectors for multi-scale multi-frequency CLEAN from THE
REFERENCE'S OWN msmfsclean() (/root/reference/src/ska_sdp_func_python/image/cleaners.py:686-898;
pure numpy, imported unmodified), run in the build container:
    python tests/golden/make_golden_msmfsclean.py"""
import os

import numpy as np

from make_golden_clean import HERE, load_cleaners, make_images
from make_golden_msclean import extended


def moment_images(seed, nx, ny, px, py, nmoment, nsrc):
    """Frequency-moment dirty images and PSFs of a sky whose sources have spectral slopes: moment
    t of the dirty image = sum over 5 channels of w_c^t x (channel image), PSF moments likewise
    (2 nmoment of them)."""
    rng = np.random.default_rng(seed)
    nchan = 5
    wc = np.linspace(-0.2, 0.2, nchan)                      # (nu - nu0) / nu0
    _, psf0 = make_images(seed, nx, ny, px, py, 1)
    yy, xx = np.meshgrid(np.arange(py) - py // 2, np.arange(px) - px // 2)
    r = np.hypot(xx, yy)
    chan_psf = [psf0 * (1.0 + 0.3 * w * np.cos(r / 3.0) * (r > 0)) for w in wc]
    flux = rng.uniform(0.5, 3.0, nsrc)
    slope = rng.uniform(-1.0, 1.0, nsrc)
    sx, sy = rng.integers(4, nx - 4, nsrc), rng.integers(4, ny - 4, nsrc)
    chans = []
    for c in range(nchan):
        model = np.zeros((nx, ny))
        model[sx, sy] = flux * (1.0 + slope * wc[c])
        model = extended(model, seed + 1) if c == 0 else model + (extended(np.zeros((nx, ny)), seed + 1))
        big = np.zeros((nx, ny))
        cx, cy = min(px, nx), min(py, ny)
        big[:cx, :cy] = chan_psf[c][px // 2 - cx // 2: px // 2 - cx // 2 + cx, py // 2 - cy // 2: py // 2 - cy // 2 + cy]
        big = np.roll(big, (-(cx // 2), -(cy // 2)), axis=(0, 1))
        chans.append(np.real(np.fft.ifft2(np.fft.fft2(model) * np.fft.fft2(big))) + 0.01 * rng.normal(size=(nx, ny)))
    dirty = np.array([sum(wc[c] ** t * chans[c] for c in range(nchan)) / nchan for t in range(nmoment)])
    psf = np.array([sum(wc[c] ** t * chan_psf[c] for c in range(nchan)) / nchan for t in range(2 * nmoment)])
    return dirty, psf / psf[0].max()


def main():
    ref = load_cleaners()
    out = {}
    cases = [
        # name, nx, ny, px, py, nmoment, window, sensitivity, gain, thresh, niter, fracthresh, scales
        # (cases e, f: findpeak="CASA")
        ("a", 64, 64, 64, 64, 3, None, False, 0.7, 0.0, 50, 0.01, [0, 3, 10]),
        ("b", 48, 72, 32, 40, 2, "quarter", False, 0.3, 0.0, 80, 0.02, [0, 4]),
        ("c", 64, 64, 64, 64, 2, "float", True, 0.5, 0.0, 40, 0.05, [0, 3, 8]),
        ("d", 40, 40, 40, 40, 1, None, False, 0.7, 1.2, 400, 0.0, [0, 3]),       # one moment; threshold stop
        ("e", 64, 64, 64, 64, 3, None, False, 0.7, 0.0, 40, 0.01, [0, 3, 10]),
        ("f", 48, 72, 32, 40, 2, "quarter", True, 0.3, 0.0, 60, 0.02, [0, 4]),
    ]
    for name, nx, ny, px, py, nm, wk, use_sens, gain, thresh, niter, frac, scales in cases:
        dirty, psf = moment_images(90 + ord(name), nx, ny, px, py, nm, 8)
        if nm == 1:
            psf = psf[:2]
        if wk is None:
            window = None
        elif wk == "quarter":
            window = np.zeros((nx, ny), dtype=bool)
            window[nx // 4: 3 * nx // 4, ny // 4: 3 * ny // 4] = True
        else:
            window = (np.random.default_rng(3).uniform(0.0, 1.0, (nx, ny)) > 0.02).astype(float)
        sens = None
        if use_sens:
            xx, yy = np.meshgrid(np.linspace(-1, 1, nx), np.linspace(-1, 1, ny), indexing="ij")
            sens = np.exp(-(xx ** 2 + yy ** 2))
        findpeak = "CASA" if name in "ef" else "RASCIL"
        model, res = ref.msmfsclean(dirty.copy(), psf.copy(), window, sens, gain, thresh, niter, scales, frac,
                                    findpeak=findpeak)
        out.update({f"{name}_dirty": dirty, f"{name}_psf": psf, f"{name}_model": model,
                    f"{name}_res": res, f"{name}_par": np.array([gain, thresh, niter, frac]),
                    f"{name}_scales": np.array(scales, dtype=float)})
        if window is not None:
            out[f"{name}_window"] = window
        if sens is not None:
            out[f"{name}_sens"] = sens
        print(name, "nonzero model", int((model != 0).sum()), "res max", np.abs(res).max(axis=(1, 2)))
    np.savez_compressed(os.path.join(HERE, "msmfsclean_case.npz"), **out)


if __name__ == "__main__":
    main()
