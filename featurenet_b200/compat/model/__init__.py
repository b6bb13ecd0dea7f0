"""Architecture description of a FeatureNet candidate (block / cell / input / operation / output).

Same module and class names as /root/reference/model/, so the reference drivers and the
mutation operators find what they import; the layers they instantiate are the symbolic ones
of ``featurenet_b200.layers`` and the resulting ``Model`` is the CUDA executor.
"""
