"""Drop-in replacement for the reference file lib/ransac_voting_gpu_layer/ransac_voting_gpu.py.

Copy this file over the reference's module (and put `pvnet_b200/` on PYTHONPATH).  The callers
keep their imports unchanged:
    from lib.ransac_voting_gpu_layer.ransac_voting_gpu import ransac_voting_layer_v3, \\
        estimate_voting_distribution_with_mean                      (tools/train_linemod.py:21-22, tools/demo.py:8)
Re-exported: the two functions on the B200 path, the wrappers built on the same kernels (ransac_voting_layer_v5,
generate_hypothesis, ransac_motion_voting), plus the primitive-compatible module object under the reference's name
`ransac_voting` for the wrappers that still call generate_hypothesis / voting_for_hypothesis.
"""
from pvnet_b200.ransac_voting_gpu import (estimate_voting_distribution_with_mean,  # noqa: F401
                                          generate_hypothesis, ransac_motion_voting, ransac_voting,
                                          ransac_voting_layer_v3, ransac_voting_layer_v5)
