"""Importable alias for the package directory `clustermachinelearningforhospitalnetworks-apache-spark_b200/`
(its mandated name is not a Python identifier).  `import hkm_b200` and
`from hkm_b200.ml.clustering import KMeans` both work and share module objects with the real package."""
import importlib
import sys

_REAL = "clustermachinelearningforhospitalnetworks-apache-spark_b200"
_pkg = importlib.import_module(_REAL)
for _name, _mod in list(sys.modules.items()):
    if _name == _REAL or _name.startswith(_REAL + "."):
        sys.modules["hkm_b200" + _name[len(_REAL):]] = _mod
sys.modules[__name__] = _pkg
