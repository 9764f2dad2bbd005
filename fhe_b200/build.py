"""Build libluxb200.so in-tree with nvcc for sm_100a (no torch extension machinery: the product boundary is a
plain C ABI).  `python -m fhe_b200.build` or fhe_b200.build.build()."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libluxb200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr", "-Xptxas", "-v"]
OBJ = os.path.join(HERE, "_obj")
# translation units: api.cu holds every kernel of the bootstrapping / NTT path (minutes to compile); circuit.cu is the
# level scheduler above it
UNITS = ["api.cu", "br11.cu", "circuit.cu", "rns.cu"]


def sources():
    return [os.path.join(CSRC, u) for u in UNITS]


def _newest_dep(unit):
    """api.cu includes every other file of csrc/; circuit.cu only the headers."""
    t = os.path.getmtime(os.path.join(HERE, "..", "include", "luxb200.h"))
    for f in os.listdir(CSRC):
        if f.endswith(".cu") and f != unit:
            continue
        if unit != "api.cu" and f.endswith(".inc"):
            continue
        t = max(t, os.path.getmtime(os.path.join(CSRC, f)))
    return t


def _obj(unit):
    return os.path.join(OBJ, unit.replace(".cu", ".o"))


def _stale(unit):
    o = _obj(unit)
    return not os.path.exists(o) or os.path.getmtime(o) < _newest_dep(unit)


def needs_build():
    if not os.path.exists(OUT):
        return True
    return any(_stale(u) for u in UNITS) or any(os.path.getmtime(_obj(u)) > os.path.getmtime(OUT) for u in UNITS)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    os.makedirs(OBJ, exist_ok=True)
    log = os.path.join(HERE, "build.log")
    text = ""
    procs = []
    for u in UNITS:
        if force or _stale(u):
            # br11.cu is 16 large kernels: its optimisation passes run on all host cores (api.cu keeps the single-threaded
            # pipeline its tuned kernels were measured with)
            extra = ["--split-compile", "0"] if u == "br11.cu" else []
            cmd = [NVCC] + FLAGS + extra + ["-c", "-o", _obj(u), os.path.join(CSRC, u)]
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for cmd, p in procs:
        out = p.communicate()[0]
        text += " ".join(cmd) + "\n" + out
        failed |= p.returncode != 0
    if not failed:
        cmd = [NVCC, "-shared", "-o", OUT] + [_obj(u) for u in UNITS]
        r = subprocess.run(cmd, capture_output=True, text=True)
        text += " ".join(cmd) + "\n" + r.stdout + r.stderr
        failed = r.returncode != 0
    with open(log, "w") as f:
        f.write(text)
    if failed:
        sys.stderr.write(text)
        raise RuntimeError("nvcc failed; see " + log)
    if verbose:
        print(text)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print("built", OUT)
