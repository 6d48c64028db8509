#!/usr/bin/env python
"""Golden vectors for the complex (Q + iU) Hogbom CLEAN from THE REFERENCE'S OWN hogbom_complex()
(/root/reference/src/ska_sdp_func_python/image/cleaners.py:136-232; pure numpy, imported
unmodified), run in the build container:   python tests/golden/make_golden_clean_complex.py"""
import os

import numpy as np

from make_golden_clean import HERE, load_cleaners, make_images


def main():
    ref = load_cleaners()
    out = {}
    cases = [
        # name, nx, ny, px, py, window, gain, thresh, niter, fracthresh, psf scale
        ("a", 64, 64, 64, 64, None, 0.1, 0.0, 250, 0.01, 1.0),
        ("b", 48, 80, 33, 41, "quarter", 0.25, 0.0, 300, 0.05, 1.0),        # window; stops at the threshold
        ("c", 40, 40, 80, 80, "float", 0.7, 0.0, 60, 0.0, 1.7),             # PSF peak != 1: Smith's division
        ("d", 32, 32, 32, 32, None, 0.2, 0.0, 6, 0.0, 1.0),
    ]
    for name, nx, ny, px, py, wk, gain, thresh, niter, frac, pscale in cases:
        dq, psf = make_images(70 + ord(name), nx, ny, px, py, 10)
        du, _ = make_images(170 + ord(name), nx, ny, px, py, 10)
        psf = psf * pscale
        if wk is None:
            window = None
        elif wk == "quarter":
            window = np.zeros((nx, ny), dtype=bool)
            window[nx // 4: 3 * nx // 4, ny // 4: 3 * ny // 4] = True
        else:
            window = np.random.default_rng(3).uniform(0.0, 1.0, (nx, ny))
        cq, cu, rq, ru = ref.hogbom_complex(dq, du, psf, psf.copy(), window, gain, thresh, niter, frac)
        out.update({f"{name}_dq": dq, f"{name}_du": du, f"{name}_psf": psf, f"{name}_cq": cq,
                    f"{name}_cu": cu, f"{name}_rq": rq, f"{name}_ru": ru,
                    f"{name}_par": np.array([gain, thresh, niter, frac])})
        if window is not None:
            out[f"{name}_window"] = window
        print(name, "components", int(((cq != 0) | (cu != 0)).sum()), "res max", np.hypot(rq, ru).max())
    np.savez_compressed(os.path.join(HERE, "clean_complex_case.npz"), **out)


if __name__ == "__main__":
    main()
