"""Compile the C oracle (test infrastructure) into oracle/_build/liblcoracle.so with gcc."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, '_build')
LIB = os.path.join(OUT, 'liblcoracle.so')
SRC = os.path.join(HERE, 'lc_oracle.c')


def build(force=False):
    os.makedirs(OUT, exist_ok=True)
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    cmd = ['gcc', '-O2', '-fopenmp', '-ffp-contract=off', '-fno-fast-math', '-mf16c', '-shared', '-fPIC',
           SRC, '-o', LIB, '-lm']
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f'gcc failed:\n{r.stdout}\n{r.stderr}')
    return LIB


if __name__ == '__main__':
    print(build(force=True))
