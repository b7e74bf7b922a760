"""In-tree build of the CUDA library (sm_100a only).  `python -m ktx_software_b200.build` or build_ext()."""
import os, subprocess, sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libuastc_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "--fmad=false", "-std=c++17",
         "-Xcompiler", "-fPIC", "-shared"]


def _sources():
    out = []
    for d, _, fs in os.walk(CSRC):
        out += [os.path.join(d, f) for f in fs if f.endswith((".cu", ".cuh", ".inc", ".h"))]
    out.append(os.path.join(os.path.dirname(HERE), "include", "ktx_b200_uastc.h"))
    return out


def build_id():
    """First 12 hex digits of the SHA-1 over every source the library is compiled from: names the build in bench lines and in
    profiles/ncu_counters.json, so a counter measured on one build is never quoted for another."""
    import hashlib
    h = hashlib.sha1()
    for f in sorted(_sources()):
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:12]


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in _sources())


def build_ext(force=False, verbose=False):
    if not force and not stale():
        return LIB
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ['-DB200_BUILD_ID="%s"' % build_id()] + ["-o", LIB, os.path.join(CSRC, "uastc_b200.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode:
        raise RuntimeError("nvcc failed: " + " ".join(cmd))
    return LIB


if __name__ == "__main__":
    print(build_ext(force="--force" in sys.argv, verbose=True))
