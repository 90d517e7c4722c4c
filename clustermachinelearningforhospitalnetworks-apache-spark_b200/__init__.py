"""B200-native KMeans hot path behind the pyspark.ml Estimator/Transformer surface."""
