"""Puts the REFERENCE's pure-Python `torchrec` package under baseline/_ref/ (git-ignored; it travels to the GPU box)
so that tests can run the reference's own DistributedModelParallel / ShardedEmbeddingBagCollection /
BatchedFusedEmbeddingBag over THIS repo's `fbgemm_gpu` (the drop-in claim of INTEGRATION.md, Option A).

The sanctioned `pip install --no-index --no-build-isolation --target baseline/_ref /root/reference` cannot work:
the reference's setup.py first builds third_party/fbgemm/fbgemm_gpu with CUDA
(`setup.py build -DTORCH_CUDA_ARCH_LIST=7.0;8.0`, needs cmake + a scikit-build tree) and fails at metadata
generation.  torchrec itself is pure Python, so the package directory is copied instead - unmodified - together
with two five-line stand-ins for its import-time dependencies that are not in this image (iopath, pyre_extensions;
SURVEY.md section 8c).  Nothing is written inside the repository's tracked tree.

usage: python tools/install_reference_torchrec.py [/root/reference]
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main() -> int:
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    src = os.path.join(ref, "torchrec")
    if not os.path.isdir(src):
        print(f"no reference checkout at {ref}: nothing installed")
        return 0
    dst_root = os.path.join(ROOT, "baseline", "_ref")
    dst = os.path.join(dst_root, "torchrec")
    shutil.rmtree(dst, ignore_errors=True)
    shutil.copytree(src, dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    stubs = {
        "iopath/__init__.py": "",
        "iopath/common/__init__.py": "",
        "iopath/common/file_io.py":
            "class PathManager:\n    def open(self, path, mode='r', **kw):\n        return open(path, mode)\n\n\n"
            "class PathManagerFactory:\n    @staticmethod\n    def get(key=None, **kw):\n        return PathManager()\n",
        "pyre_extensions/__init__.py":
            "import typing\n\nParameterSpecification = typing.ParamSpec\n\n\n"
            "def none_throws(x, msg='Unexpected None'):\n    assert x is not None, msg\n    return x\n",
    }
    for rel, text in stubs.items():
        path = os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "w") as f:
            f.write(text)
    print(f"reference torchrec -> {dst} (+ iopath / pyre_extensions stand-ins)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
