#!/usr/bin/env python
"""Golden vectors for multi-scale CLEAN from THE REFERENCE'S OWN
msclean() (/root/reference/src/ska_sdp_func_python/image/cleaners.py:279-470; pure numpy, imported
unmodified), run in the build container:   python tests/golden/make_golden_msclean.py"""
import os

import numpy as np

from make_golden_clean import HERE, load_cleaners, make_images


def extended(dirty, seed):
    """Add a few broad blobs so that the larger scales get picked."""
    rng = np.random.default_rng(seed)
    nx, ny = dirty.shape
    xx, yy = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    for _ in range(3):
        cx, cy = rng.uniform(0.25, 0.75) * nx, rng.uniform(0.25, 0.75) * ny
        dirty = dirty + rng.uniform(1.0, 2.5) * np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * rng.uniform(2.0, 5.0) ** 2))
    return dirty


def main():
    ref = load_cleaners()
    out = {}
    cases = [
        # name, nx, ny, px, py, window, sensitivity, gain, thresh, niter, fracthresh, scales
        ("a", 64, 64, 64, 64, None, False, 0.7, 0.0, 60, 0.01, [0, 3, 10]),
        ("b", 48, 80, 32, 40, "quarter", False, 0.25, 0.05, 150, 0.0, [0, 4]),
        ("c", 64, 64, 64, 64, "float", True, 0.5, 0.0, 40, 0.05, [0, 3, 8, 14]),
        ("d", 32, 32, 32, 32, None, False, 0.1, 0.0, 7, 0.0, [0]),
        ("e", 56, 40, 56, 40, None, False, 0.7, 0.6, 500, 0.0, [0, 3, 6]),   # stops at the threshold
    ]
    for name, nx, ny, px, py, wk, use_sens, gain, thresh, niter, frac, scales in cases:
        dirty, psf = make_images(60 + ord(name), nx, ny, px, py, 8)
        dirty = extended(dirty, 7 + ord(name))
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
        comps, res = ref.msclean(dirty, psf, window, sens, gain, thresh, niter, scales, frac)
        out.update({f"{name}_dirty": dirty, f"{name}_psf": psf, f"{name}_comps": comps,
                    f"{name}_res": res, f"{name}_par": np.array([gain, thresh, niter, frac]),
                    f"{name}_scales": np.array(scales, dtype=float)})
        if window is not None:
            out[f"{name}_window"] = window
        if sens is not None:
            out[f"{name}_sens"] = sens
        print(name, "nonzero comps", int((comps != 0).sum()), "res max", np.abs(res).max())
    np.savez_compressed(os.path.join(HERE, "msclean_case.npz"), **out)


if __name__ == "__main__":
    main()
