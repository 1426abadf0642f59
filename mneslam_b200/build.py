"""Builds ``libmneslam_b200.so`` in-tree with nvcc for sm_100a (no torch C++ headers involved: the library
is a plain C-ABI shared object, loaded with ctypes)."""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB_PATH = os.path.join(HERE, 'libmneslam_b200.so')
OBJ_DIR = os.path.join(os.path.dirname(HERE), 'build', 'obj')
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC', '-Xptxas', '-v']


def _nvcc():
    exe = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(exe):
        raise RuntimeError('nvcc not found: cannot build libmneslam_b200.so')
    return exe


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith('.cu'))


def _deps_mtime():
    paths = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    paths.append(os.path.join(os.path.dirname(HERE), 'include', 'mneslam_b200.h'))
    return max(os.path.getmtime(p) for p in paths)


def needs_build():
    return (not os.path.exists(LIB_PATH)) or os.path.getmtime(LIB_PATH) < _deps_mtime()


def build(force=False, verbose=True):
    """Compile every .cu under csrc/ (in parallel) and link the shared library.  Returns its path."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = _nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    dep_time = _deps_mtime()

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, src[:-3] + '.o')
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= dep_time:
            return obj, ''
        cmd = [nvcc] + NVCC_FLAGS + ['-c', os.path.join(CSRC, src), '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed for %s:\n%s\n%s' % (src, r.stdout, r.stderr))
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=8) as ex:
        results = list(ex.map(compile_one, sources()))
    objs = [o for o, _ in results]
    if verbose:
        for _, log in results:
            for line in log.splitlines():
                if 'registers' in line or 'spill' in line and 'bytes spill' in line and ' 0 bytes spill stores' not in line:
                    print('[mneslam_b200.build]', line.strip(), file=sys.stderr)
    r = subprocess.run([nvcc, '-shared', '-o', LIB_PATH] + objs + ['-lcudart'], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n%s\n%s' % (r.stdout, r.stderr))
    return LIB_PATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
