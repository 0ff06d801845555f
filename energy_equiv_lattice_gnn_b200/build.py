"""Build libeeg_b200.so (sm_100a only) in-tree with nvcc; no torch headers are involved (plain C ABI)."""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libeeg_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "--use_fast_math" if os.environ.get("EEG_FAST_MATH") else "-DEEG_PRECISE"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps(path: str, seen=None):
    """The file itself + every quoted #include it reaches (recursively)."""
    import re
    seen = set() if seen is None else seen
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    with open(path) as f:
        for inc in re.findall(r'^\s*#\s*include\s+"([^"]+)"', f.read(), flags=re.M):
            _deps(os.path.join(os.path.dirname(path), inc), seen)
    return seen


def _digest(src: str) -> str:
    h = hashlib.sha256()
    for p in sorted(_deps(src)):
        with open(p, "rb") as f:
            h.update(p.encode() + b"\0" + f.read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Incremental: an object is recompiled only when its source, a header it includes, or the flags changed."""
    bdir = os.path.join(HERE, "build")
    os.makedirs(bdir, exist_ok=True)
    srcs = sources()

    def cc(src):
        base = os.path.basename(src)[:-3]
        obj, stamp = os.path.join(bdir, base + ".o"), os.path.join(bdir, base + ".stamp")
        dig = _digest(src)
        if not force and not verbose and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
            return obj, False
        cmd = [NVCC, *FLAGS, "-c", src, "-o", obj] + (["-Xptxas", "-v"] if verbose else [])
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        with open(stamp, "w") as f:
            f.write(dig)
        return obj, True

    with cf.ThreadPoolExecutor(max_workers=8) as ex:
        res = list(ex.map(cc, srcs))
    objs = [o for o, _ in res]
    link_stamp = os.path.join(bdir, "link.stamp")
    link_key = "\n".join(sorted(objs))
    if any(ch for _, ch in res) or not os.path.exists(LIB) or not os.path.exists(link_stamp) or open(link_stamp).read() != link_key:
        r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs, "-lcudart"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        with open(link_stamp, "w") as f:
            f.write(link_key)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
