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


# ---- multi-scale CLEAN (cleaners.py:279-470 and helpers :473-610, :613-683) -------------------
# Pinned against tests/golden/msclean_case.npz (outputs of the reference's own msclean(),
# tests/golden/make_golden_msclean.py) in tests/test_oracle_clean.py.

_PSWF_P = numpy.array([[8.203343e-2, -3.644705e-1, 6.278660e-1, -5.335581e-1, 2.312756e-1],
                       [4.028559e-3, -3.697768e-2, 1.021332e-1, -1.201436e-1, 6.412774e-2]])
_PSWF_Q = numpy.array([[1.0, 8.212018e-1, 2.078043e-1], [1.0, 9.599102e-1, 2.918724e-1]])


def pswf(nu):
    """Schwab's rational approximation of the m=6, alpha=1 prolate spheroidal wave function
    (spheroidal_function, cleaners.py:613-683) for one scalar nu; 0 outside [0, 1]."""
    if not 0.0 <= nu <= 1.0:
        return 0.0
    part, nuend = (0, 0.75) if nu < 0.75 else (1, 1.0)
    d = nu ** 2 - nuend ** 2
    top = _PSWF_P[part, 0]
    for k in range(1, 5):
        top += _PSWF_P[part, k] * d ** k
    bot = _PSWF_Q[part, 0]
    for k in range(1, 3):
        bot += _PSWF_Q[part, k] * d ** k
    return max(top / bot, 0.0) if bot != 0.0 else 0.0


def scale_stack(shape, scales, norm=True):
    """create_scalestack (cleaners.py:473-508): tapered-paraboloid blobs of the given widths, each
    normalised to unit sum; width 0 is a delta.  Centre = ceil(n / 2)."""
    ns, nx, ny = shape
    assert ns == len(scales)
    out = numpy.zeros(shape)
    xc, yc = int(numpy.ceil(nx / 2.0)), int(numpy.ceil(ny / 2.0))
    for s, width in enumerate(scales):
        if width > 0.0:
            half = int(numpy.ceil(width / 2.0))
            inv = 1.0 / (float(width) / 2.0) ** 2
            for y in range(yc - half - 1, yc + half + 1):
                for x in range(xc - half - 1, xc + half + 1):
                    fx, fy = float(x - xc), float(y - yc)
                    r = numpy.sqrt(inv * (fx * fx + fy * fy))
                    out[s, x, y] = pswf(r) * (1.0 - r ** 2)
            out[out < 0.0] = 0.0
            if norm:
                out[s] /= numpy.sum(out[s])
        else:
            out[s, xc, yc] = 1.0
    return out


def _centred_spectrum(a):
    # the reference wraps every transform in fftshift on both sides; the frequency-side shifts are
    # permutations applied to every factor of an element-wise product and cancel exactly
    return numpy.fft.fft2(numpy.fft.fftshift(a))


def convolve_stack(stack, img):
    """convolve_scalestack (cleaners.py:511-531): img correlated with every scale."""
    fi = _centred_spectrum(img)
    return numpy.stack([numpy.real(numpy.fft.ifftshift(numpy.fft.ifft2(fi * numpy.conjugate(_centred_spectrum(s)))))
                        for s in stack])


def convolve_convolve_stack(stack, img):
    """convolve_convolve_scalestack (cleaners.py:534-562): out[s, p] = img * scale_p (x) scale_s."""
    fi = _centred_spectrum(img)
    fs = [_centred_spectrum(s) for s in stack]
    ns = len(fs)
    out = numpy.zeros((ns, ns) + img.shape)
    for s in range(ns):
        for p in range(ns):
            out[s, p] = numpy.real(numpy.fft.ifftshift(numpy.fft.ifft2(fi * fs[p] * numpy.conjugate(fs[s]))))
    return out


def stack_peak(stack, sensitivity, windowstack, coupling):
    """find_max_abs_stack (cleaners.py:565-610), including its double use of the sensitivity in
    the argmax and the strict '>' across scales."""
    best, bx, by, bs = 0.0, 0, 0, 0
    for s in range(stack.shape[0]):
        r = stack[s] * windowstack[s] / coupling[s, s] if windowstack is not None else stack[s] / coupling[s, s]
        if sensitivity is not None:
            r = r * sensitivity
            k = numpy.abs(r * sensitivity).argmax()
        else:
            k = numpy.abs(r).argmax()
        x, y = numpy.unravel_index(k, r.shape)
        if numpy.abs(r[x, y]) > best:
            best, bx, by, bs = numpy.abs(r[x, y]), x, y, s
    return bx, by, bs


def msclean_setup(dirty, psf, window, scales):
    """Everything msclean() builds before its loop (cleaners.py:320-383)."""
    pmax = psf.max()
    assert pmax > 0.0
    lpsf, ldirty = psf / pmax, dirty / pmax
    ns = len(scales)
    sstack = scale_stack((ns,) + ldirty.shape, scales)
    pstack = scale_stack((ns,) + lpsf.shape, scales)
    res = convolve_stack(sstack, numpy.array(ldirty))
    pss = convolve_convolve_stack(pstack, numpy.array(lpsf))
    coupling = pss.max(axis=(2, 3))
    wstack = None
    if window is not None:
        wstack = numpy.zeros_like(sstack)
        wstack[convolve_stack(sstack, window) > 0.9] = 1.0
        assert numpy.sum(wstack) > 0
    return pmax, pstack, res, pss, coupling, wstack


def msclean_loop(res, pss, pstack, coupling, wstack, sensitivity, pshape, gain, absolutethresh, niter):
    """The minor cycle proper (cleaners.py:398-451) on prepared stacks; res is updated in place.
    Returns (comps, subtractions done)."""
    shape = res.shape[1:]
    comps = numpy.zeros(shape)
    done = 0
    for _ in range(niter):
        mx, my, ms = stack_peak(res, sensitivity, wstack, coupling)
        mval = res[ms, mx, my] / coupling[ms, ms]
        if numpy.abs(res[ms, mx, my]) < 0.9 * absolutethresh or not numpy.abs(mval) > 0:
            break
        a1, a2 = overlap(shape, pshape, mx, my)
        for s in range(res.shape[0]):
            res[s, a1[0]:a1[1], a1[2]:a1[3]] -= pss[s, ms, a2[0]:a2[1], a2[2]:a2[3]] * gain * mval
        comps[a1[0]:a1[1], a1[2]:a1[3]] += pstack[ms, a2[0]:a2[1], a2[2]:a2[3]] * gain * mval
        done += 1
    return comps, done


def msclean(dirty, psf, window, sensitivity, gain, thresh, niter, scales, fracthresh):
    """Returns (comps, residual, subtractions done)."""
    assert 0.0 < gain < 2.0 and niter > 0 and len(scales) > 0
    pmax, pstack, res, pss, coupling, wstack = msclean_setup(dirty, psf, window, scales)
    absolutethresh = max(thresh, fracthresh * numpy.max(numpy.abs(res[0])))
    comps, done = msclean_loop(res, pss, pstack, coupling, wstack, sensitivity, psf.shape, gain,
                               absolutethresh, niter)
    return comps, pmax * res[0], done


# ---- restoring beam (image/deconvolution.py:1016-1036) -------------------------------------------
# PARITY UNPINNED: the reference builds the beam with astropy.convolution.Gaussian2DKernel and
# convolves with astropy's convolve_fft(boundary="wrap"); astropy is not installed in the build
# container, so this restates their documented behaviour (kernel of odd size ceil(8 max sigma),
# sampled at pixel centres, peak-normalised; circular convolution centred on the kernel's middle).

def beam_kernel(sx, sy, theta):
    size = int(numpy.ceil(8.0 * max(sx, sy)))
    if size % 2 == 0:
        size += 1
    h = size // 2
    k = numpy.zeros((size, size))
    ct, st = numpy.cos(theta), numpy.sin(theta)
    for iy in range(size):
        for ix in range(size):
            xr = (ix - h) * ct + (iy - h) * st          # along the x_stddev axis
            yr = -(ix - h) * st + (iy - h) * ct
            k[iy, ix] = numpy.exp(-0.5 * ((xr / sx) ** 2 + (yr / sy) ** 2))
    return k / k.max()


def restore_plane(model, residual, sx, sy, theta):
    """Direct (non-FFT) circular convolution of one plane with the beam, plus the residual."""
    from scipy import ndimage
    out = ndimage.convolve(model, beam_kernel(sx, sy, theta), mode="wrap")
    return out + residual if residual is not None else out


# ---- multi-scale multi-frequency CLEAN (cleaners.py:686-898 and helpers :901-1160) ---------------
# Pinned against tests/golden/msmfsclean_case.npz (outputs of the reference's own msmfsclean(),
# tests/golden/make_golden_msmfsclean.py) in tests/test_oracle_clean.py.

def msmfs_setup(dirty, psf, window, scales):
    """Everything msmfsclean() builds before its loop (cleaners.py:752-795): scale shapes, the
    scale-convolved moment residuals [ns, nm, nx, ny], the coupled PSFs [ns, ns, nm, nm, px, py],
    the per-scale moment Hessians (PSF centre values) and their inverses, the window stack."""
    pmax = psf.max()
    assert pmax > 0.0
    lpsf, ldirty = psf / pmax, dirty / pmax
    nm, ns = dirty.shape[0], len(scales)
    if nm > 1:
        assert psf.shape[0] == 2 * nm
    sstack = scale_stack((ns,) + ldirty.shape[1:], scales)
    pstack = scale_stack((ns,) + lpsf.shape[1:], scales)
    smres = numpy.zeros((ns, nm) + ldirty.shape[1:])
    for t in range(nm):
        smres[:, t] = convolve_stack(sstack, ldirty[t])
    nmp = max(lpsf.shape[0] // 2, 1)
    ssmm = numpy.zeros((ns, ns, nmp, nmp) + lpsf.shape[1:])
    for t in range(nmp):
        for q in range(nmp):
            ssmm[:, :, t, q] = convolve_convolve_stack(pstack, lpsf[t + q])
    hess = numpy.array([ssmm[s, s, :, :, lpsf.shape[1] // 2, lpsf.shape[2] // 2] for s in range(ns)])
    ihess = numpy.array([numpy.linalg.inv(h) for h in hess])
    wstack = None
    if window is not None:
        wstack = numpy.zeros_like(sstack)
        wstack[convolve_stack(sstack, window) > 0.9] = 1.0
    return pmax, pstack, smres, ssmm, hess, ihess, wstack


def msmfs_peak(smres, ihess, wstack, sensitivity, hess=None, findpeak="RASCIL"):
    """find_global_optimum, RASCIL / Algorithm1 branch (cleaners.py:901-975, :1124-1157): the scale
    is chosen on the windowed, sensitivity-weighted principal solution of moment 0, the pixel is the
    peak of the UN-windowed one."""
    sol = numpy.einsum("smn,smxy->snxy", ihess, smres)
    search = sol
    if findpeak == "CASA":
        # CASA 4.7 (cleaners.py:936-960): search the decrease of chi^2 instead
        ns, nm = sol.shape[:2]
        search = numpy.zeros((ns, 1) + sol.shape[2:])
        for s in range(ns):
            for m1 in range(nm):
                search[s, 0] += 2.0 * sol[s, m1] * smres[s, m1]
                for m2 in range(nm):
                    search[s, 0] -= hess[s, m1, m2] * sol[s, m1] * sol[s, m2]
    best, bs, bx, by = 0.0, 0, 0, 0
    for s in range(sol.shape[0]):
        r = search[s, 0] * wstack[s] if wstack is not None else search[s, 0]
        if sensitivity is not None:
            r = sensitivity * r
        top = numpy.max(numpy.abs(r))
        if top > best:
            best, bs = top, s
            bx, by = numpy.unravel_index(numpy.abs(search[s, 0]).argmax(), search[s, 0].shape)
    return bs, bx, by, sol[bs, :, bx, by]


def msmfs_loop(smres, ssmm, pstack, ihess, wstack, sensitivity, gain, absolutethresh, niter,
               hess=None, findpeak="RASCIL"):
    """The minor cycle (cleaners.py:826-880); smres is updated in place.  Returns (model, done)."""
    ns, nm = smres.shape[:2]
    shape, pshape = smres.shape[2:], ssmm.shape[4:]
    model = numpy.zeros((nm,) + shape)
    done = 0
    for _ in range(niter):
        ms, mx, my, mval = msmfs_peak(smres, ihess, wstack, sensitivity, hess, findpeak)
        if numpy.fabs(mval[0]) < absolutethresh:
            break
        a1, a2 = overlap(shape, pshape, mx, my)
        for t in range(nm):
            model[t, a1[0]:a1[1], a1[2]:a1[3]] += pstack[ms, a2[0]:a2[1], a2[2]:a2[3]] * gain * mval[t]
        smres[:, :, a1[0]:a1[1], a1[2]:a1[3]] -= gain * numpy.einsum(
            "stqxy,q->stxy", ssmm[ms, :, :, :, a2[0]:a2[1], a2[2]:a2[3]], mval)
        done += 1
    return model, done


def msmfsclean(dirty, psf, window, sensitivity, gain, thresh, niter, scales, fracthresh, findpeak="RASCIL"):
    """dirty [nmoment, nx, ny], psf [2 nmoment, px, py].  Returns (moment model, moment residual,
    components subtracted)."""
    assert 0.0 < gain < 2.0 and niter > 0 and len(scales) > 0
    pmax, pstack, smres, ssmm, hess, ihess, wstack = msmfs_setup(dirty, psf, window, scales)
    absolutethresh = max(thresh, fracthresh * numpy.max(numpy.abs(smres[0, 0])))
    model, done = msmfs_loop(smres, ssmm, pstack, ihess, wstack, sensitivity, gain, absolutethresh, niter,
                             hess, findpeak)
    return model, pmax * smres[0], done


# ---- complex Hogbom CLEAN of Q + iU (cleaners.py:136-232) -----------------------------------------
# Pinned bit-exact against tests/golden/clean_complex_case.npz (outputs of the reference's own
# hogbom_complex(), tests/golden/make_golden_clean_complex.py).

def hogbom_complex(dirty_q, dirty_u, psf_q, psf_u, window, gain, thresh, niter, fracthresh):
    """Returns (comps_q, comps_u, res_q, res_u, iterations).  numpy complex128 arithmetic, as the
    reference: note the stop test without the 0.9 of the real-valued hogbom()."""
    assert 0.0 < gain < 2.0 and niter > 0
    assert numpy.all(psf_q == psf_u)
    z = dirty_q + 1j * dirty_u
    absolutethresh = max(thresh, fracthresh * numpy.absolute(z).max())
    comps = numpy.zeros(z.shape, dtype="complex128")
    res = numpy.array(z)
    pmax = psf_q.max()
    assert pmax > 0.0
    it = 0
    for i in range(niter):
        it = i + 1
        a = numpy.absolute(res * window) if window is not None else numpy.absolute(res)
        mx, my = numpy.unravel_index(a.argmax(), z.shape)
        mval = res[mx, my] * gain / pmax
        comps[mx, my] += mval
        a1, a2 = overlap(z.shape, psf_q.shape, mx, my)
        res[a1[0]:a1[1], a1[2]:a1[3]] -= psf_q[a2[0]:a2[1], a2[2]:a2[3]] * mval
        if numpy.abs(res[mx, my]) < absolutethresh:
            break
    return comps.real, comps.imag, res.real, res.imag, it
