"""B200-native KMeans hot path behind the pyspark.ml Estimator/Transformer surface.

    from hkm_b200.ml.feature import VectorAssembler, StandardScaler
    from hkm_b200.ml.clustering import KMeans
    from hkm_b200.ml.evaluation import ClusteringEvaluator

(`hkm_b200` is the importable alias of this directory, whose mandated name contains hyphens.)
"""
from . import _native, _runtime, _sparkrand, table  # noqa: F401
from .table import Table, DeviceMatrix, Column, AssembledColumn  # noqa: F401
from .ml import param, linalg, feature, clustering, evaluation, _kmeans_init  # noqa: F401
from .ml.feature import VectorAssembler, StandardScaler, StandardScalerModel, SparkException  # noqa: F401
from .ml.evaluation import ClusteringEvaluator  # noqa: F401
from . import datagen, io  # noqa: F401
from .ml.clustering import KMeans, KMeansModel, KMeansSummary  # noqa: F401
from .ml.param import IllegalArgumentException  # noqa: F401
from ._native import HkmError  # noqa: F401
