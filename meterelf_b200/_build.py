"""Builds ``_lib/libmeterelf_b200.so`` in-tree with nvcc for sm_100a.

The flags matter for parity: ``-fmad=false`` keeps the float32 HLS arithmetic and the
float64 angle arithmetic from being contracted into FMAs (the kernels also use explicit
``__f*_rn`` / ``__d*_rn`` intrinsics on those paths).
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT_DIR = os.environ.get('MELF_OUT_DIR') or os.path.join(HERE, '_lib')
LIB = os.path.join(OUT_DIR, 'libmeterelf_b200.so')

SOURCES = ['melf_api.cu', 'melf_match.cu', 'melf_match_tc.cu', 'melf_dials.cu', 'melf_convert.cu', 'melf_calib.cu', 'melf_jpeg.cu', 'melf_host.cpp']
NVCC_FLAGS = [
    '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
    '-fmad=false', '-Xcompiler', '-fPIC,-ffp-contract=off,-Wall', '-Xptxas', '-v',
]


def _nvcc() -> str:
    path = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(path):
        raise RuntimeError('nvcc not found')
    return path


def _stale(target: str, deps: List[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.h', '.cuh'))]
    headers.append(os.path.join(os.path.dirname(HERE), 'include', 'meterelf_b200.h'))
    nvcc = _nvcc()
    objs = []
    jobs = []
    for src in SOURCES:
        path = os.path.join(CSRC, src)
        obj = os.path.join(OUT_DIR, os.path.splitext(src)[0] + '.o')
        objs.append(obj)
        if force or _stale(obj, [path] + headers):
            jobs.append([nvcc] + NVCC_FLAGS + os.environ.get('MELF_EXTRA_NVCC_FLAGS', '').split() +
                        ['-c', path, '-o', obj])

    def run(cmd):
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        return (cmd, res.returncode, res.stdout)
    with ThreadPoolExecutor(max_workers=4) as pool:
        for (cmd, rc, out) in pool.map(run, jobs):
            if verbose or rc:
                sys.stderr.write(' '.join(cmd) + '\n' + out)
            with open(cmd[-1] + '.log', 'w') as fp:
                fp.write(out)
            if rc:
                raise RuntimeError('nvcc failed for ' + cmd[-3])
    if force or jobs or _stale(LIB, objs):
        cmd = [nvcc, '-shared', '-o', LIB] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a']
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if res.returncode:
            sys.stderr.write(res.stdout)
            raise RuntimeError('link failed')
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose=True))
