"""Relink libhkm.so from the objects under csrc/build (after compiling single units by hand)."""
import importlib, os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
B = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._build")
objs = [os.path.join(B.OBJ, n + ".o") for n, _, _ in B._units(False)]
r = subprocess.run(["nvcc", "-shared", "-o", B.SO, *objs, "-gencode", "arch=compute_100a,code=sm_100a"], capture_output=True, text=True)
print("relink", r.returncode, r.stderr[-300:])
sys.exit(r.returncode)
