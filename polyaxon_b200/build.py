"""In-tree build of libplxbo.so (hand-written CUDA for sm_100a + the C ABI).  nvcc cross-compiles
without a GPU; the .so is git-ignored but travels to the GPU box with the repo snapshot."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT = os.path.join(HERE, 'libplxbo.so')
OUT_PROFILE = os.path.join(HERE, 'libplxbo_profile.so')     # separate file: never loaded by the product
SOURCES = ['api.cu']
HEADERS = ['mma_f64.cuh', 'tma_ring.cuh', 'gp_kernels.cuh', 'sweep.cuh', 'sweep_i8.cuh', 'sweep_i8f.cuh', 'sweep_i8c.cuh', 'sweep_i8p.cuh', 'mt19937.cuh', 'mt19937_jump.inc', 'fit_small.cuh', 'sweep_small.cuh', 'append.cuh', 'refine.cuh', 'lbfgsb1d.h', os.path.join('..', '..', 'include', 'plxbo.h')]
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-shared', '-Xcompiler', '-fPIC']


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return all(os.path.getmtime(p) <= t for p in deps if os.path.exists(p))


def build(force=False, verbose=False, profile=False):
    """profile=True adds -DPLX_PROFILE: the work-skipping `sweep_dbg` switches that tools/*_phases.py use to time
    parts of the sweep kernels.  They are compiled out of the normal build."""
    if not force and not profile and up_to_date():
        return OUT
    out = OUT_PROFILE if profile else OUT
    nvcc = os.environ.get('NVCC', 'nvcc')
    cmd = [nvcc] + NVCC_FLAGS + (['-DPLX_PROFILE'] if profile else []) + (['-Xptxas', '-v'] if verbose else []) + ['-o', out] + SOURCES
    res = subprocess.run(cmd, cwd=CSRC, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError('nvcc failed building libplxbo.so')
    if verbose:
        print(res.stdout)
    return out


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv, profile='--profile' in sys.argv))
