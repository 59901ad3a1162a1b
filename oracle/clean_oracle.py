"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's Hogbom CLEAN,
the parity oracle of sdp_hogbom_clean. Never imported by the product.

Follows the reference's host path (/root/reference/src/ska-sdp-func/):
  clean/sdp_hogbom_clean.cpp:33-80    sdp_create_cbeam      -> clean_beam
  clean/sdp_hogbom_clean.cpp:183-231  the minor cycle       -> minor_cycles
  numeric_functions/sdp_fft_convolution.cpp:127-244 (pad both images to a power
      of two >= n + m - 1 at (P - size) / 2, FFT, multiply, inverse FFT, swap
      halves, crop from (P - n) / 2 - 1)                    -> fft_convolution
  clean/sdp_hogbom_clean.cpp:244-262  restore step          -> hogbom_clean
The minor cycle keeps the host loop's arithmetic: the peak is the FIRST largest
element (np.argmax), `highest_value` is a double, the update is
residual - (loop_gain * highest_value) * psf with every product rounded (no
fused multiply-add) and, for fp32 images, evaluated in double and rounded once
to fp32. Pinned by tests/golden/golden_clean.npz, outputs of the unmodified
reference library's CPU path (tests/golden/make_golden_clean.py): model and
residual bit for bit, skymodel to FFT rounding.
"""
import numpy as np


def clean_beam(details, real=np.float64):
    """.cpp:33-80; evaluated in double, cast to the image type."""
    sx, sy, theta_deg, size = (float(v) for v in details)
    dim = int(size)
    theta = (np.pi / 180) * theta_deg
    a = np.cos(theta) ** 2 / (2 * sx ** 2) + np.sin(theta) ** 2 / (2 * sy ** 2)
    b = np.sin(2 * theta) / (4 * sx ** 2) - np.sin(2 * theta) / (4 * sy ** 2)
    c = np.sin(theta) ** 2 / (2 * sx ** 2) + np.cos(theta) ** 2 / (2 * sy ** 2)
    d = np.arange(dim, dtype=np.float64) - dim // 2
    dx, dy = d[:, None], d[None, :]
    return np.exp(-(a * dx * dx + 2 * b * dx * dy + c * dy * dy)).astype(real)


def minor_cycles(dirty, psf, loop_gain, threshold, cycle_limit):
    """-> (clean_model, residual, cycles). .cpp:183-231."""
    real = dirty.dtype.type
    n = dirty.shape[0]
    residual = dirty.copy()
    model = np.zeros_like(dirty)
    gain = real(loop_gain)
    thr = real(threshold)
    cycles = 0
    while cycles < cycle_limit:
        flat = int(np.argmax(residual))          # first of equal maxima
        mx, my = divmod(flat, n)
        peak = np.float64(residual[mx, my])
        if residual[mx, my] < thr:
            break
        scale = np.float64(gain) * peak           # loop_gain * highest_value
        model[mx, my] += real(scale)
        window = psf[n - mx:2 * n - mx, n - my:2 * n - my]
        residual = (residual.astype(np.float64)
                    - scale * window.astype(np.float64)).astype(real)
        cycles += 1
    return model, residual, cycles


def fft_convolution(in1, in2):
    """sdp_fft_convolution.cpp:127-244 for square real images. The reference
    is only sound for EVEN sizes: its padding takes `cols + 2 * pad` for the
    row stride of the padded image (:66), which is P - 1 when P - size is
    odd (a skewed beam; for an odd image size the crop writes (n + 1)^2
    elements into n^2: heap corruption, seen as a crash). Here the padded
    image is always P x P, i.e. the even-size rule for every size."""
    n, m = in1.shape[0], in2.shape[0]
    p = n + m - 1
    while p & (p - 1):
        p += 1
    e1, e2 = (p - n) // 2, (p - m) // 2
    a = np.zeros((p, p), dtype=np.complex128)
    b = np.zeros((p, p), dtype=np.complex128)
    a[e1:e1 + n, e1:e1 + n] = in1
    b[e2:e2 + m, e2:e2 + m] = in2
    c = np.fft.ifft2(np.fft.fft2(a) * np.fft.fft2(b))
    half = p // 2
    c = np.roll(c, (half, half), axis=(0, 1))    # swap halves (p is even)
    return c[e1 - 1:e1 - 1 + n, e1 - 1:e1 - 1 + n].real


def hogbom_clean(dirty, psf, cbeam_details, loop_gain, threshold, cycle_limit):
    """-> dict(clean_model, residual, skymodel, cycles)."""
    real = dirty.dtype.type
    model, residual, cycles = minor_cycles(dirty, psf, loop_gain, threshold,
                                           cycle_limit)
    beam = clean_beam(cbeam_details, real)
    sky = fft_convolution(model.astype(np.float64), beam.astype(np.float64))
    return dict(clean_model=model, residual=residual,
                skymodel=(sky.astype(real) + residual), cycles=cycles)
