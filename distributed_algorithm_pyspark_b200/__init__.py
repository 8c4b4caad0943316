"""distributed_algorithm_pyspark_b200 -- a B200-native (sm_100a) distributed-algorithms framework with
the capabilities of orion-orion/Distributed-Algorithm-PySpark (PageRank, transitive closure, k-means,
logistic regression, ALS / Gram matrix decomposition, SSGD / MA / BMUF / EASGD, Monte-Carlo pi).

Layers:
  oracle/    NumPy restatement with exact Spark semantics (numerical ground truth, CPU baseline)
  parallel/  one-process-per-GPU runtime, logical partitions, symmetric heap over NVLink peer memory
  data/      reference datasets + synthetic generators for the benchmark configurations
  models/    the algorithms, each with ``oracle`` / ``torch`` (NCCL baseline) / ``fused`` backends
  ops/       thin Python wrappers over the C ABI of the native kernels (include/dapb200.h)
  utils/     Philox twin, timing, plotting, reporting
"""
__version__ = "0.1.0"

from .config import (ALSConfig, KMeansConfig, LogisticConfig, MonteCarloConfig, PageRankConfig,
                     TransitiveClosureConfig)

__all__ = ["ALSConfig", "KMeansConfig", "LogisticConfig", "MonteCarloConfig", "PageRankConfig",
           "TransitiveClosureConfig", "__version__"]
