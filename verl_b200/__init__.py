"""verl_b200 — B200-native (sm_100a) implementation of verl's per-token policy-gradient hot path.

Same Python API as the reference for that path, hand-written CUDA behind a C ABI underneath
(include/vpg_b200.h, verl_b200/csrc).  There is no Triton, no multi-backend dispatch and no CPU fallback.

    from verl_b200.utils.kernel.linear_cross_entropy import linear_cross_entropy
    from verl_b200.utils import torch_functional as verl_F
    from verl_b200.trainer.ppo import core_algos
    from verl_b200 import bert_padding          # unpad_input / pad_input / index_first_axis
    import verl_b200.compat; verl_b200.compat.install()   # shadow the same names inside an installed verl
"""

__version__ = "0.1.0"
