"""Regenerates tests/golden/*.json.  Run from the repo root: python tests/golden/make_golden.py

Sources of the constants (none come from /root/reference, which holds no KMeans code or tests):
  * java.util.Random outputs            -- JDK specification (LCG), well-known values
  * pyspark rand(seed=42)/randn(seed=42) doc examples (pyspark.sql.functions.rand / randn,
    Spark 3.5 API docs): XORShiftRandom(42 + partitionIndex 0) first two draws
  * iris silhouettes 0.6564679231 / 0.7222369298 -- Spark ClusteringEvaluatorSuite constants
    (relTol 1e-5), iris in the UCI variant (rows 35 and 38 differ from scikit-learn's copy)
  * KMeansSuite "single cluster": points (1,2,6),(1,3,0),(1,4,6), k=1 -> centre (1,3,4)
  * pyspark.ml.clustering.KMeans doctest: 4 weighted points, k=2 -> trainingCost 4.0
  * pyspark.ml.evaluation.ClusteringEvaluator doctest (Spark 3.5 API docs, pyspark.ml.evaluation.ClusteringEvaluator): six points
    ([0.0,0.5],0) ([0.5,0.0],0) ([10.0,11.0],1) ([10.5,11.5],1) ([1.0,1.0],0) ([8.0,6.0],1) -> evaluate() prints 0.9079...
    The doctest then sets weightCol (weights 2.5 x4, 5.0 x2) and prints 0.9079... again in some doc versions; applying
    ClusteringMetrics' published formulas to those weights gives 0.87153 (hand-checked point by point), so only the
    UNWEIGHTED constant is treated as an upstream known answer; the weighted value recorded here is the formula's
  * pyspark.ml.feature.StandardScaler doctest (API docs): rows [0.0], [2.0] -> mean [1.0], std [1.4142], transform -> [0.0], [1.4142]
  * pyspark.ml.feature.VectorAssembler doctest (API docs): row (a=1, b=0, c=3) -> features [1.0, 0.0, 3.0]
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
        "clustering_evaluator_doctest": {"points": [[0.0, 0.5], [0.5, 0.0], [10.0, 11.0], [10.5, 11.5], [1.0, 1.0], [8.0, 6.0]],
                                         "prediction": [0, 0, 1, 1, 0, 1], "silhouette_prefix": 0.9079, "prefix_tol": 5e-5,
                                         "weights": [2.5, 2.5, 2.5, 2.5, 5.0, 5.0], "weighted_by_formula": 0.8715279275368312},
        "standard_scaler_doctest": {"rows": [[0.0], [2.0]], "mean": [1.0], "std": [1.4142], "scaled": [[0.0], [1.4142]], "print_tol": 5e-5},
        "vector_assembler_doctest": {"a": [1], "b": [0], "c": [3], "features": [[1.0, 0.0, 3.0]]},
    }
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(kat, f, indent=1)
    np.savetxt(os.path.join(HERE, "iris_uci.csv"), np.column_stack([X, y]), fmt="%.1f", delimiter=",")


if __name__ == "__main__":
    main()
