"""CPU oracle (TEST INFRASTRUCTURE ONLY) for the Hogbom row N4: numpy restatement of hogbom() and
overlapIndices() (/root/reference/src/ska_sdp_func_python/image/cleaners.py:23-133, :235-267).
Pinned bit-exact against tests/golden/clean_case.npz, which holds outputs of the reference function
itself (tests/golden/make_golden_clean.py; tests/test_oracle_clean.py)."""
import numpy


def overlap(shape, pshape, peakx, peaky):
    nx, ny = shape
    pwx, pwy = pshape[0] // 2, pshape[1] // 2
    rl = (max(0, peakx - pwx), max(0, peaky - pwy))
    ru = (min(nx, peakx + pwx), min(peaky + pwy, ny))
    pl = (max(0, pwx + (rl[0] - peakx)), max(0, pwy + (rl[1] - peaky)))
    pu = (min(pshape[0], pwx + (ru[0] - peakx)), min(pwy + (ru[1] - peaky), pshape[1]))
    return (rl[0], ru[0], rl[1], ru[1]), (pl[0], pu[0], pl[1], pu[1])


def hogbom(dirty, psf, window, gain, thresh, niter, fracthresh):
    """Returns (comps, res, iterations)."""
    assert 0.0 < gain < 2.0 and niter > 0
    absolutethresh = max(thresh, fracthresh * numpy.fabs(dirty).max())
    comps = numpy.zeros(dirty.shape)
    res = numpy.array(dirty)
    pmax = psf.max()
    it = 0
    for i in range(niter):
        it = i + 1
        a = numpy.fabs(res * window) if window is not None else numpy.fabs(res)
        mx, my = numpy.unravel_index(a.argmax(), dirty.shape)
        mval = res[mx, my] * gain / pmax
        comps[mx, my] += mval
        a1, a2 = overlap(dirty.shape, psf.shape, mx, my)
        res[a1[0]:a1[1], a1[2]:a1[3]] -= psf[a2[0]:a2[1], a2[2]:a2[3]] * mval
        if numpy.abs(res[mx, my]) < 0.9 * absolutethresh:
            break
    return comps, res, it
