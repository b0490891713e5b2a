"""TEST INFRASTRUCTURE - builds oracle/c/wn_oracle.c into oracle/_build/libwn_oracle.so with gcc.

-ffp-contract=off keeps gcc from fusing a*b+c on its own (the explicit fmaf calls are the only fused operations,
mirroring the explicit __fmaf_rn calls of the CUDA kernels); -mfma makes fmaf a single hardware instruction.
There is no reference C/C++ source to compile (the reference is 100% Python/TensorFlow), so there is
no oracle/_ref."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'c', 'wn_oracle.c')
OUT_DIR = os.path.join(HERE, '_build')
OUT = os.path.join(OUT_DIR, 'libwn_oracle.so')


def build(force=False):
    deps = [SRC, os.path.join(HERE, 'c', 'wn_oracle.h'), os.path.join(HERE, '..', 'include', 'wn_math.h')]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        return OUT
    os.makedirs(OUT_DIR, exist_ok=True)
    cmd = ['gcc', '-O3', '-std=c11', '-mavx2', '-fno-math-errno', '-fPIC', '-shared', '-ffp-contract=off', '-mfma', '-fopenmp',
           '-fno-fast-math', SRC, '-o', OUT, '-lm']
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
