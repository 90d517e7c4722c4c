"""In-tree nvcc build of csrc/ -> csrc/libhkm.so (sm_100a only).

The .so is git-ignored but travels to the GPU box with the repo snapshot; `build()` is what
__graft_entry__.build() calls.  Translation units compile in parallel; unchanged objects are
reused (mtime check against every header).
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
SO = os.path.join(CSRC, "libhkm.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
]


def _units(dev_fast: bool):
    units = [("hkm_api", "hkm_api.cu", [])]
    for extra in sorted(glob.glob(os.path.join(CSRC, "hkm_*.cu"))):
        base = os.path.basename(extra)
        if base in ("hkm_api.cu", "hkm_lloyd_inst.cu", "hkm_lloyd_tp_inst.cu"):
            continue
        units.append((base[:-3], base, []))
    for nt in range(1, 9):
        flags = [f"-DHKM_NT={nt}"] + (["-DHKM_DEV_FAST"] if dev_fast else [])
        units.append((f"hkm_lloyd_nt{nt}" + ("_dev" if dev_fast else ""), "hkm_lloyd_inst.cu", flags))
    for kp in (16, 32, 48, 64):
        units.append((f"hkm_lloyd_tp{kp}", "hkm_lloyd_tp_inst.cu", [f"-DHKM_TP_KP={kp}"]))
    return units


def _newest_header() -> float:
    hs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(_HERE, "..", "include", "*.h"))
    return max(os.path.getmtime(h) for h in hs)


def build(force: bool = False, dev_fast: bool | None = None, verbose: bool = False) -> str:
    if dev_fast is None:
        dev_fast = os.environ.get("HKM_DEV_FAST", "0") == "1"
    os.makedirs(OBJ, exist_ok=True)
    hdr = _newest_header()
    jobs = []
    objs = []
    for name, src, flags in _units(dev_fast):
        srcp = os.path.join(CSRC, src)
        obj = os.path.join(OBJ, name + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(srcp), hdr):
            cmd = ["nvcc", *NVCC_FLAGS, *flags, "-c", srcp, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
            logs = list(ex.map(run, jobs))
        if verbose:
            sys.stderr.write("\n".join(logs))
    if jobs or not os.path.exists(SO):
        cmd = ["nvcc", "-shared", "-o", SO, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
