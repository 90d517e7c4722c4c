"""Regenerates tests/golden/*.json.  Run from the repo root: python tests/golden/make_golden.py

Sources of the constants (none come from /root/reference, which holds no KMeans code or tests):
  * java.util.Random outputs            -- JDK specification (LCG), well-known values
  * pyspark rand(seed=42)/randn(seed=42) doc examples (pyspark.sql.functions.rand / randn,
    Spark 3.5 API docs): XORShiftRandom(42 + partitionIndex 0) first two draws
  * iris silhouettes 0.6564679231 / 0.7222369298 -- Spark ClusteringEvaluatorSuite constants
    (relTol 1e-5), iris in the UCI variant (rows 35 and 38 differ from scikit-learn's copy)
  * KMeansSuite "single cluster": points (1,2,6),(1,3,0),(1,4,6), k=1 -> centre (1,3,4)
  * pyspark.ml.clustering.KMeans doctest: 4 weighted points, k=2 -> trainingCost 4.0
The iris matrix itself is written by this script from scikit-learn's copy with the two UCI rows patched.
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    from sklearn.datasets import load_iris
    X, y = load_iris(return_X_y=True)
    X = X.copy()
    X[34] = [4.9, 3.1, 1.5, 0.1]
    X[37] = [4.9, 3.1, 1.5, 0.1]
    kat = {
        "java_random": {"seed0_nextInt": -1155484576, "seed42_nextInt": -1170105035, "seed0_nextDouble": 0.730967787376657,
                        "seed42_nextInt10_x5": [0, 3, 8, 4, 0]},
        "xorshift": {"seed42_rand_times3": [1.8575681106759028, 1.5288056527339444],
                     "seed42_randn": [2.384479054241165, 0.1920934041293524]},
        "iris": {"silhouette_sqeuclidean": 0.6564679231, "silhouette_cosine": 0.7222369298, "rel_tol": 1e-5},
        "kmeans_single_cluster": {"points": [[1, 2, 6], [1, 3, 0], [1, 4, 6]], "center": [1, 3, 4]},
        "pyspark_doctest": {"points": [[0, 0], [1, 1], [9, 8], [8, 9]], "weights": [2, 2, 2, 2], "k": 2, "trainingCost": 4.0,
                            "centers_sorted": [[0.5, 0.5], [8.5, 8.5]]},
        "kmeans_default_seed": -1689246527,
    }
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(kat, f, indent=1)
    np.savetxt(os.path.join(HERE, "iris_uci.csv"), np.column_stack([X, y]), fmt="%.1f", delimiter=",")


if __name__ == "__main__":
    main()
