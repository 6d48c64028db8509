#!/usr/bin/env python
"""Golden vectors for the Hogbom row from THE REFERENCE'S OWN hogbom()
(/root/reference/src/ska_sdp_func_python/image/cleaners.py:23-133; pure numpy, imported
unmodified), run in the build container:   python tests/golden/make_golden_clean.py"""
import importlib.util
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def load_cleaners():
    spec = importlib.util.spec_from_file_location(
        "ref_cleaners", "/root/reference/src/ska_sdp_func_python/image/cleaners.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def make_images(seed, nx, ny, px, py, nsrc):
    """A PSF with unit peak and broad sidelobes, and a dirty image = sources convolved with it
    (circularly) plus noise."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(py) - py // 2, np.arange(px) - px // 2)
    r = np.hypot(xx, yy)
    psf = np.exp(-(r / 1.7) ** 2) + 0.08 * np.cos(r / 2.3) * np.exp(-r / 18.0) * (r > 0)
    psf[px // 2, py // 2] = 1.0
    model = np.zeros((nx, ny))
    model[rng.integers(0, nx, nsrc), rng.integers(0, ny, nsrc)] = rng.uniform(0.2, 3.0, nsrc) * \
        rng.choice([-1.0, 1.0], nsrc, p=[0.2, 0.8])
    big = np.zeros((nx, ny))
    sx, sy = min(px, nx), min(py, ny)
    big[:sx, :sy] = psf[px // 2 - sx // 2: px // 2 - sx // 2 + sx, py // 2 - sy // 2: py // 2 - sy // 2 + sy]
    big = np.roll(big, (-(sx // 2), -(sy // 2)), axis=(0, 1))
    dirty = np.real(np.fft.ifft2(np.fft.fft2(model) * np.fft.fft2(big))) + 0.01 * rng.normal(size=(nx, ny))
    return dirty, psf


def main():
    ref = load_cleaners()
    out = {}
    cases = [
        # name, nx, ny, px, py, window kind, gain, thresh, niter, fracthresh
        ("a", 64, 64, 64, 64, None, 0.1, 0.0, 300, 0.01),
        ("b", 48, 80, 33, 41, "quarter", 0.25, 0.05, 500, 0.0),       # odd, smaller PSF; window
        ("c", 40, 40, 80, 80, "float", 0.7, 0.0, 120, 0.2),           # PSF larger than the image
        ("d", 32, 32, 32, 32, None, 0.1, 0.0, 7, 0.0),                # niter < 10
    ]
    for name, nx, ny, px, py, wk, gain, thresh, niter, frac in cases:
        dirty, psf = make_images(40 + ord(name), nx, ny, px, py, 12)
        if wk is None:
            window = None
        elif wk == "quarter":
            window = np.zeros((nx, ny), dtype=bool)
            window[nx // 4: 3 * nx // 4, ny // 4: 3 * ny // 4] = True
        else:
            window = np.random.default_rng(3).uniform(0.0, 1.0, (nx, ny))
        comps, res = ref.hogbom(dirty, psf, window, gain, thresh, niter, frac)
        out.update({f"{name}_dirty": dirty, f"{name}_psf": psf, f"{name}_comps": comps,
                    f"{name}_res": res, f"{name}_par": np.array([gain, thresh, niter, frac])})
        if window is not None:
            out[f"{name}_window"] = window
        print(name, "components", int((comps != 0).sum()), "res max", np.abs(res).max())
    np.savez_compressed(os.path.join(HERE, "clean_case.npz"), **out)


if __name__ == "__main__":
    main()
