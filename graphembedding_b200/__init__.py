"""graphembedding_b200: B200-native (sm_100a) Siamese graph-pair similarity path of
kangzf/GraphEmbedding (model/Siamese): hand-written CUDA behind a C ABI (include/sgnn.h),
with a Python host side mirroring the reference's flags, layer-spec grammar, data loaders and
similarity / metrics interfaces."""
__version__ = '0.1.0'
