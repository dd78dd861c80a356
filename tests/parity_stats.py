"""Statistics shared by the converged-render parity tests.

z scores of an image against a reference whose per-sample variance is known, under the null
hypothesis that both sample the same per-pixel distribution (so one variance serves both):
    se^2 = var (1/n_img + 1/n_ref)
Pixels are split in two classes:
  * flat pixels — relative standard error below 1e-4 (a constant background, the inside of a
    uniformly lit object): there is no measured noise to normalise by (4 096 f32 additions of the same
    value carry a coherent rounding error of a few 1e-6, which a z score would blow up into a "bias"),
    so they are compared directly: flat_max_rel = the largest |difference| as a fraction of the frame's
    mean radiance.  It is not zero — a pixel looks flat when none of the reference's samples met the
    object whose rim crosses 1/5000 of it, and the other render's one sample that did moves the pixel by
    ~1e-3 — but a wrong background or a shifted flat region is 10 to 1000 times that;
  * noisy pixels: per pixel and channel the fraction of |z| > 4 (outliers: a feature one side renders
    and the other does not), and per 4x4-pixel block the mean and rms of z.  Bias is judged on blocks,
    not on pixels: a path-traced pixel is a skewed distribution (rare bright paths), its sample mean
    sits below the true mean most of the time and its sample variance knows nothing of the events it
    missed, so pixel-level z of a better-converged image against it has a positive mean by
    construction (measured: +0.16 on the book-3 fuzz scenes with identical frame means).  Sixteen
    pixels per block restore the symmetry and make the test four times more sensitive to a uniform bias.
"""
import numpy as np


def z_stats(img, n_img, ref, var, n_ref, block=4):
    se2 = var * (1.0 / n_img + 1.0 / n_ref)
    finite = np.isfinite(img) & np.isfinite(ref)
    noisy = finite & (se2 > (1e-4 * np.abs(ref)) ** 2)
    flat = finite & ~noisy
    flat_rel = np.abs(img[flat] - ref[flat]) / float(np.mean(np.abs(ref[finite])))
    z = (img[noisy] - ref[noisy]) / np.sqrt(se2[noisy])
    H, W = img.shape[0] // block * block, img.shape[1] // block * block
    sh = (H // block, block, W // block, block, img.shape[2])
    d = np.where(noisy, img - ref, 0.0)[:H, :W].reshape(sh).sum(axis=(1, 3))
    s = np.where(noisy, se2, 0.0)[:H, :W].reshape(sh).sum(axis=(1, 3))
    zb = d[s > 0] / np.sqrt(s[s > 0])
    return dict(block_mean_z=float(zb.mean()), block_rms_z=float(np.sqrt((zb * zb).mean())), blocks=int(zb.size),
                pixel_mean_z=float(np.clip(z, -6, 6).mean()), frac_over_4=float((np.abs(z) > 4).mean()),
                max_abs_z=float(np.abs(z).max()), n=int(z.size), flat=int(flat.sum()),
                flat_max_rel=float(flat_rel.max()) if flat_rel.size else 0.0)
