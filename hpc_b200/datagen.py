"""Fixed-seed synthetic inputs with the reference's distributions.

cux fills arrays with `rand() % mod + min` integer-valued floats seeded by
time(0) (0-frameworks/cux/include/util/data_filler.h:41-48); the samples use
`rand() % 20 - 10` (cuda/gemm_fp16_wmma.cu:46) and `i%9 + j%9`
(simd/x86/linear.cpp:19-25).  We keep the distributions and replace the seed
with a constant so the CPU oracle and the GPU see identical bytes.
"""
import numpy as np

SEED = 20240607


def rng(tag=0):
    return np.random.default_rng(SEED + int(tag))


def int_fill(shape, lo, hi, dtype=np.float32, tag=0):
    """Integers in [lo, hi] (inclusive, like DataFiller::RandomFill with decimal_pose=0)."""
    return rng(tag).integers(lo, hi + 1, size=shape).astype(dtype)


def uniform_fill(shape, lo=-1.0, hi=1.0, tag=0):
    return rng(tag).uniform(lo, hi, size=shape).astype(np.float32)


def normal_fill(shape, tag=0):
    return rng(tag).standard_normal(size=shape).astype(np.float32)


def linear_sample_fill(rows, cols, dtype=np.float32):
    """GenMatrix of simd/x86/linear.cpp:19-25: mat[i,j] = i%9 + j%9 (values 0..16)."""
    i = np.arange(rows)[:, None] % 9
    j = np.arange(cols)[None, :] % 9
    return (i + j).astype(dtype)
