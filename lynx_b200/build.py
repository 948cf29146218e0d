"""Build ``lib/liblynx_b200.so`` in-tree with nvcc for sm_100a.

    python -m lynx_b200.build [--force]

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels with the tree.
"""

import os
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
SRC = ["api.cu", "compose.cu", "track.cu", "track_pipe.cu", "adjoint.cu", "host.cu", "exchange.cu", "generate.cu", "beamops.cu", "along.cu", "noalloc.cu"]
OUT = HERE / "lib" / "liblynx_b200.so"
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "--threads", "0",
]


def needs_build() -> bool:
    if not OUT.exists():
        return True
    t = OUT.stat().st_mtime
    deps = list((HERE / "csrc").glob("*")) + [ROOT / "include" / "lynx_b200.h"]
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False, experiments: bool = False) -> Path:
    """``experiments``: a second library, ``liblynx_b200_exp.so``, compiled with
    -DLYNX_EXPERIMENTS (tuning knobs read from the environment; select it with LYNX_B200_LIB).
    The product library reads no environment variable."""
    out = OUT.with_name("liblynx_b200_exp.so") if experiments else OUT
    if not force and not experiments and not needs_build():
        return OUT
    OUT.parent.mkdir(parents=True, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, *FLAGS, f"-I{ROOT / 'include'}", "-o", str(out)]
    if experiments:
        cmd += ["-DLYNX_EXPERIMENTS"]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [str(HERE / "csrc" / s) for s in SRC]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv,
                experiments="--experiments" in sys.argv))
