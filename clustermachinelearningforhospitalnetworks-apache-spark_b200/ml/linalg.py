"""Minimal pyspark.ml.linalg stand-in: dense vectors are NumPy fp64 arrays."""
import numpy as np


class Vectors:
    @staticmethod
    def dense(*values):
        if len(values) == 1 and not np.isscalar(values[0]):
            values = values[0]
        return np.asarray(values, dtype=np.float64)
