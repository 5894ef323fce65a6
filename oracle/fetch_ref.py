"""Test infrastructure, not product code: place the UNMODIFIED python reference where it can travel to the GPU box.

The reference (blufzzz/learnable-triangulation-pytorch) is pure python; `/root/reference` only exists in the build
container. This recipe copies its `mvn` package (models + utils, ~3 k lines, no data) verbatim into `oracle/_ref/mvn`:
  * `oracle/_ref/` is listed in .gitignore (the sources never enter this repo's history) but not in .gpurunignore, so
    the copy ships with the gpurun snapshot like a built .so;
  * only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of `bench.py` may import
    it (through `ref_path()`); nothing under `learnable-triangulation-pytorch_b200/` does.
Used by tests/test_gpu_reference_nets.py (the reference's own `VolumetricTriangulationNet` / `AlgebraicTriangulationNet`
forward + backward on the B200, stock vs. patched with the drop-ins) and by `bench.py --impl reference`
(`cpu_baseline.kind == "reference"`).

    python oracle/fetch_ref.py            # no-op with a message when /root/reference is absent
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("LT_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")


def ref_path():
    """Directory to put on sys.path so that `import mvn` finds the unmodified reference, or None."""
    if os.path.isdir(os.path.join(SRC, "mvn")):
        return SRC
    if os.path.isdir(os.path.join(DST, "mvn")):
        return DST
    return None


def fetch(verbose=True):
    src = os.path.join(SRC, "mvn")
    if not os.path.isdir(src):
        if verbose:
            print("fetch_ref: %s not present; keeping %s as it is" % (src, DST))
        return os.path.isdir(os.path.join(DST, "mvn"))
    dst = os.path.join(DST, "mvn")
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    shutil.copytree(src, dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc", "human36m_preprocessing"))
    if verbose:
        n = sum(len(f) for _, _, f in os.walk(dst))
        print("fetch_ref: copied %d files of the unmodified reference to %s" % (n, dst))
    return True


if __name__ == "__main__":
    sys.exit(0 if fetch() else 1)
