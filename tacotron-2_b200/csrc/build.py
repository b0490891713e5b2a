"""Builds libwavenet_b200.so (hand-written sm_100a CUDA + the C ABI of include/wavenet_b200.h) with nvcc, in-tree.

--fmad=false: nvcc never fuses a*b+c on its own; the kernels write every fma explicitly (__fmaf_rn), which keeps
their rounding identical to the CPU checker.  -lineinfo maps ncu source pages to these files."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, '..', 'wavenet_vocoder', 'libwavenet_b200.so')
SOURCES = ['wn_api.cu', 'wn_upsample.cu', 'wn_loop_generic.cu', 'wn_loop_pipe.cu', 'wn_loop_tc.cu', 'wn_loop_tc2.cu', 'wn_step_full.cu', 'wn_step_gemm.cu', 'wn_step_umma.cu']


def build(force=False, verbose=False):
    out = os.path.abspath(OUT)
    srcs = [os.path.join(HERE, s) for s in SOURCES]
    deps = srcs + [os.path.join(HERE, 'wn_common.cuh'), os.path.join(HERE, '..', '..', 'include', 'wavenet_b200.h'),
                   os.path.join(HERE, '..', '..', 'include', 'wn_math.h')]
    deps += [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith('.cuh')]
    if not force and os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
        return out
    cmd = ['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17', '--fmad=false',
           '-Xcompiler', '-fPIC', '-shared', '-cudart', 'static'] + (['-Xptxas', '-v'] if verbose else []) + \
          srcs + ['-o', out]
    subprocess.run(cmd, check=True)
    return out


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
