"""pyspark.ml-shaped namespace: ml.feature, ml.clustering, ml.evaluation, ml.linalg."""
