"""alink_b200: B200-native (sm_100a) drop-in for the Lloyd-iteration hot path of
Alink's KMeansTrainBatchOp / KMeans / KMeansPredictBatchOp.

Layout: csrc/ = CUDA kernels + the C-ABI (include/akm.h); native.py = ctypes
binding; kmeans.py = host-side mirror of the reference operator API.
"""
__version__ = "0.1.0"
