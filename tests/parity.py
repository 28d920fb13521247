"""The ONE definition of the north_star pass criteria used by every GPU parity test.

fp64 mode ("within 1e-12 relative per sample"): the reference's own compiled and interpreted paths do not meet a naive
per-sample relative 1e-12 on near-zero samples, so relative means relative to signal level:

    |gpu - ref| <= 1e-12 * max(|ref[i]|, peak of the 512-frame host block holding i, 1e-3 * peak of the stream-channel)

on the doubles before the float store.  fp32 mode: RMS of (gpu float output - reference double output) below -100 dB re
full scale 1.0 over the whole render.  Integer / index state: exact."""
import numpy as np

TOL64 = 1e-12
FP32_DBFS = -100.0


def block_peak(ref, block=512):
    ns, nc, nf = ref.shape
    nb = (nf + block - 1) // block
    pad = np.zeros((ns, nc, nb * block))
    pad[:, :, :nf] = np.abs(ref)
    pk = pad.reshape(ns, nc, nb, block).max(axis=3)
    return np.repeat(pk, block, axis=2)[:, :, :nf]


def rel_err(got, ref):
    floor = np.maximum(block_peak(ref), 1e-3 * np.abs(ref).max(axis=2, keepdims=True))
    return np.abs(got - ref) / np.maximum(np.maximum(np.abs(ref), floor), 1e-300)


def dbfs_rms(r):
    return 20 * np.log10(np.sqrt(np.mean(np.asarray(r, np.float64) ** 2)) + 1e-300)
