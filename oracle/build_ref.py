#!/usr/bin/env python
"""ORACLE — TEST INFRASTRUCTURE ONLY.  Compiles the reference's OWN hot-path modules, from the sources where
they lie under /root/reference, into byte code under oracle/_ref/pyref/ (git-ignored; travels to the GPU box
with the other built artefacts):

    /root/reference/model/{model,transforms,blocks,entropy_models}.py  ->  oracle/_ref/pyref/model.*.bc
    /root/reference/utils.py                                            ->  oracle/_ref/pyref/utils.bc

The reference is pure Python, so "compiling it from its own few source files" is py_compile; no source text
is copied into the repository and nothing is modified.  The byte-code files (standard .pyc content under a
neutral extension: the snapshot that travels to the GPU box drops *.pyc) are imported through the small
meta-path finder below by tests/test_gpu_dropin.py, which runs the reference's ColorModel
UNMODIFIED over this repo's `pcc_b200.me` / `pcc_b200.compressai_compat` modules on the GPU — the drop-in
claim of INTEGRATION.md "Option A" — and by nothing else.  /root/reference does not exist on the GPU box;
__graft_entry__.build() calls build() here, in the builder container, where it does.
"""
import importlib.abc
import importlib.util
import marshal
import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref", "pyref")
FILES = ["model/model.py", "model/transforms.py", "model/blocks.py", "model/entropy_models.py", "utils.py"]


def _bc_path(rel):
    return os.path.join(OUT, rel[:-3].replace("/", ".") + ".bc")


def build(reference_root="/root/reference", force=False):
    """-> OUT when the byte code is there (built now or earlier), None when the reference is absent."""
    srcs = [os.path.join(reference_root, f) for f in FILES]
    if not all(os.path.exists(s) for s in srcs):
        return OUT if available() else None
    os.makedirs(OUT, exist_ok=True)
    for rel, src in zip(FILES, srcs):
        dst = _bc_path(rel)
        if force or not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            py_compile.compile(src, cfile=dst, dfile="reference/" + rel, doraise=True,
                               invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
    with open(os.path.join(OUT, "BUILT_WITH"), "w") as f:
        f.write(f"python {sys.version_info[0]}.{sys.version_info[1]} magic {py_compile.importlib.util.MAGIC_NUMBER.hex()}\n")
    return OUT


def available():
    """The byte code is present and was made by this interpreter's byte-code version."""
    try:
        tag = open(os.path.join(OUT, "BUILT_WITH")).read()
    except OSError:
        return False
    return importlib.util.MAGIC_NUMBER.hex() in tag and all(os.path.exists(_bc_path(rel)) for rel in FILES)


class ReferenceFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Imports `model` (a namespace-like package), `model.model|transforms|blocks|entropy_models` and `utils`
    from the byte code under OUT.  Put in front of sys.meta_path by the drop-in test, removed afterwards."""

    def __init__(self):
        self.names = {rel[:-3].replace("/", "."): rel for rel in FILES}

    def find_spec(self, name, path=None, target=None):
        if name == "model":
            return importlib.util.spec_from_loader(name, self, is_package=True)
        if name in self.names:
            return importlib.util.spec_from_loader(name, self, origin=_bc_path(self.names[name]))
        return None

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        name = module.__name__
        if name == "model":
            module.__path__ = []
            return
        path = _bc_path(self.names[name])
        with open(path, "rb") as f:
            data = f.read()
        if data[:4] != importlib.util.MAGIC_NUMBER:
            raise ImportError(f"{path} was compiled by another Python version")
        module.__file__ = path
        exec(marshal.loads(data[16:]), module.__dict__)  # 16-byte .pyc header: magic, flags, hash / mtime+size


if __name__ == "__main__":
    print(build(force="-f" in sys.argv))
