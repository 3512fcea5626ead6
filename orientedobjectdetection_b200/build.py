"""Builds librotgeo.so in-tree with nvcc for sm_100a (cross-compiles without a GPU), and the
small CPython helper _hostpack (gcc; host-side row gathering of merge_results).

    python -m orientedobjectdetection_b200.build [--force] [--verbose]
"""
import os
import subprocess
import sys
import sysconfig

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, 'csrc')
SO = os.path.join(PKG, 'librotgeo.so')
SOURCES = ['rotgeo_abi.cu', 'rotgeo_pack.cu', 'rotgeo_iou.cu', 'rotgeo_nms.cu', 'rotgeo_sort.cu', 'rotgeo_frontend.cu',
           'rotgeo_static.cu', 'rotgeo_dist.cu', 'rotgeo_assign.cu', 'rotgeo_eval.cu', 'rotgeo_diff.cu']
HOSTPACK_SRC = os.path.join(CSRC, 'hostpack.c')
HOSTPACK_SO = os.path.join(PKG, '_hostpack' + (sysconfig.get_config_var('EXT_SUFFIX') or '.so'))
HEADERS = ['rotgeo_device.cuh', 'rotgeo_internal.h', os.path.join('..', '..', 'include', 'rotgeo.h')]


def nvcc_path():
    for p in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc'):
        if p and os.path.exists(p):
            return p
    return 'nvcc'


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build_hostpack(force=False):
    """the CPython extension next to the package (plain C, the interpreter's own headers)"""
    if (not force and os.path.exists(HOSTPACK_SO)
            and os.path.getmtime(HOSTPACK_SO) >= max(os.path.getmtime(HOSTPACK_SRC), os.path.getmtime(__file__))):
        return HOSTPACK_SO
    cc = '/usr/bin/gcc' if os.path.exists('/usr/bin/gcc') else 'gcc'
    cmd = [cc, '-O2', '-Wall', '-fopenmp', '-fPIC', '-shared', '-I', sysconfig.get_paths()['include'], HOSTPACK_SRC, '-o', HOSTPACK_SO]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError('gcc failed building _hostpack')
    return HOSTPACK_SO


def build(force=False, verbose=False):
    build_hostpack(force)
    if not force and not stale():
        return SO
    cmd = [nvcc_path(), '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo',
           '-std=c++17', '-fmad=true',
           '-Xcompiler', '-fPIC,-O2,-Wall', '-shared', '-o', SO]
    if verbose:
        cmd += ['-Xptxas', '-v']
    # the image exports CC=/opt/gcc/bin/gcc (a wrapper); nvcc wants a plain host g++
    if os.path.exists('/usr/bin/g++'):
        cmd += ['-ccbin', '/usr/bin/g++']
    cmd += [os.path.join(CSRC, f) for f in SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError('nvcc failed building librotgeo.so')
    if verbose:
        sys.stderr.write(r.stdout + r.stderr)
    return SO


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='--verbose' in sys.argv))
