"""Assemble oracle/_ref/ = the UNMODIFIED reference package, so that it travels to the GPU box with the snapshot.

TEST INFRASTRUCTURE ONLY. Nothing under lightcompress_b200/ and nothing inside a timed region imports oracle/_ref;
`tests/`, `__graft_entry__` and `bench.py --impl reference` are the only users (as the checker / the baseline).

What this does (run here, where /root/reference exists; `__graft_entry__.build()` calls it):
  * oracle/_ref/llmc/            byte-for-byte copy of /root/reference/llmc (python files only), EXCEPT
        llmc/models/__init__.py  -> emptied: the original imports every model family (librosa, diffusers, ... are not
                                    installed); the harness imports llmc.models.{base_model,llama,opt} directly.
  * oracle/_ref/cpu/llmc/        the same tree after the reference's OWN CPU rewrite (ci_check/change_files.py, applied
                                    in memory: '.cuda()' -> '.to("cpu")' and friends) plus smoothquant.py:43 `x.cuda()`
                                    -> no-op, which that script misses; AWQ's n_grid is kept at 20 (the CI sets 1).
                                    Used in this container to generate tests/golden/e2e_golden.pt.gz.
  * oracle/_ref/shims/           stand-ins for two absent third-party packages the reference imports but the path never
                                    calls: `accelerate.init_empty_weights` and `easydict.EasyDict`.
  * oracle/_ref/MANIFEST.json    sha256 of every copied file against its source (tests check the copy is unmodified).
oracle/_ref/ is git-ignored (never in history) and not gpurun-ignored (ships like the built .so files).
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get('LCB_REFERENCE', '/root/reference')
OUT = os.path.join(HERE, '_ref')

# ci_check/change_files.py restated as data: (relative file, header lines, [(search, replace)]) — the reference's own
# CPU rewrite; the table below is what that script applies to the files on this path (eval / __main__ are not used).
_CPU_EDITS = {
    'compression/quantization/awq.py': (['n_grid_zbl = 20\n'],
                                        [('n_grid = 20', 'n_grid = n_grid_zbl'), ("device='cuda'", "device='cpu'")]),
    'compression/quantization/gptq.py': (['device_zbl = "cpu"\n', 'use_cuda = (device_zbl != "cpu")\n'],
                                         [('torch.cuda.empty_cache()', 'if use_cuda: torch.cuda.empty_cache()'),
                                          ('.cuda()', '.to(device_zbl)'), ("torch.device('cuda')", "torch.device('cpu')"),
                                          ('torch.cuda.synchronize()', 'if use_cuda: torch.cuda.synchronize()')]),
    'compression/quantization/base_blockwise_quantization.py': (
        ['device_zbl = "cpu"\n', 'use_cuda = (device_zbl != "cpu")\n'],
        [('.cuda()', '.to(device_zbl)'), ('torch.cuda.empty_cache()', 'if use_cuda: torch.cuda.empty_cache()')]),
    'models/base_model.py': (['device_zbl = "cpu"\n', 'use_cuda = (device_zbl != "cpu")\n'],
                             [('.cuda()', '.to(device_zbl)'),
                              ("self.move_embed_to_device('cuda')", 'self.move_embed_to_device(device_zbl)'),
                              ('torch.cuda.empty_cache()', 'if use_cuda: torch.cuda.empty_cache()')]),
    'compression/quantization/smoothquant.py': ([], [('x = x.cuda()', 'x = x')]),
    'compression/quantization/auto_clip.py': ([], [('.cuda()', '.to("cpu")'),
                                                   ('torch.cuda.empty_cache()', 'pass')]),
    'compression/quantization/module_utils.py': ([], [('torch.cuda.empty_cache()', 'pass')]),
}

_SHIMS = {
    'accelerate/__init__.py': (
        '"""Stand-in for the absent `accelerate` package: base_model.py imports init_empty_weights at module level;\n'
        'the FP8-checkpoint loader that calls it is not on the tested path."""\n'
        'import contextlib\n\n\n@contextlib.contextmanager\ndef init_empty_weights(*a, **k):\n'
        "    raise RuntimeError('accelerate is not installed (oracle/_ref shim)')\n    yield\n"),
    'easydict/__init__.py': (
        '"""Stand-in for the absent `easydict` package: attribute-accessible dict, as llmc/__main__.py builds its config."""\n\n\n'
        'class EasyDict(dict):\n'
        '    def __init__(self, d=None, **kw):\n'
        '        super().__init__()\n'
        '        for k, v in dict(d or {}, **kw).items():\n'
        '            self[k] = v\n\n'
        '    def __setitem__(self, k, v):\n'
        '        if isinstance(v, dict) and not isinstance(v, EasyDict):\n'
        '            v = EasyDict(v)\n'
        '        elif isinstance(v, (list, tuple)):\n'
        '            v = type(v)(EasyDict(x) if isinstance(x, dict) and not isinstance(x, EasyDict) else x for x in v)\n'
        '        super().__setitem__(k, v)\n\n'
        '    __setattr__ = __setitem__\n\n'
        '    def __getattr__(self, k):\n'
        '        try:\n            return self[k]\n        except KeyError:\n            raise AttributeError(k)\n\n'
        '    def update(self, *a, **kw):\n'
        '        for k, v in dict(*a, **kw).items():\n            self[k] = v\n'),
}


def _sha(path):
    with open(path, 'rb') as f:
        return hashlib.sha256(f.read()).hexdigest()


def available():
    return os.path.isdir(os.path.join(OUT, 'llmc')) and os.path.exists(os.path.join(OUT, 'MANIFEST.json'))


def build(force=False):
    """Returns the _ref path, or None when neither /root/reference nor a previous assembly exists."""
    src = os.path.join(REF_SRC, 'llmc')
    if not os.path.isdir(src):
        return OUT if available() else None
    if available() and not force:
        return OUT
    if os.path.isdir(OUT):
        shutil.rmtree(OUT)
    manifest = {}
    for variant in ('', 'cpu'):
        dst_root = os.path.join(OUT, variant, 'llmc')
        for dirpath, dirnames, filenames in os.walk(src):
            dirnames[:] = [d for d in dirnames if d != '__pycache__']
            rel_dir = os.path.relpath(dirpath, src)
            os.makedirs(os.path.join(dst_root, rel_dir), exist_ok=True)
            for fn in filenames:
                if not fn.endswith('.py'):
                    continue
                rel = os.path.normpath(os.path.join(rel_dir, fn))
                s, d = os.path.join(dirpath, fn), os.path.join(dst_root, rel)
                if rel == os.path.join('models', '__init__.py'):
                    with open(d, 'w') as f:
                        f.write('# emptied by oracle/build_ref.py: the original imports every model family\n')
                    continue
                if variant == 'cpu' and rel in _CPU_EDITS:
                    header, subs = _CPU_EDITS[rel]
                    lines = []
                    with open(s) as f:
                        for line in f:
                            for a, b in subs:
                                if a in line:
                                    line = line.replace(a, b)
                            lines.append(line)
                    with open(d, 'w') as f:
                        f.writelines(header + lines)
                    continue
                shutil.copyfile(s, d)
                if variant == '':
                    manifest[rel] = _sha(s)
    for rel, text in _SHIMS.items():
        p = os.path.join(OUT, 'shims', rel)
        os.makedirs(os.path.dirname(p), exist_ok=True)
        with open(p, 'w') as f:
            f.write(text)
    with open(os.path.join(OUT, 'MANIFEST.json'), 'w') as f:
        json.dump({'source': REF_SRC, 'files': manifest}, f, indent=0, sort_keys=True)
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
