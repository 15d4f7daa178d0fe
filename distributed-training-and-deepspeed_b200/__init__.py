"""B200-native distributed training framework (DDP, ZeRO-0..3, layer-split MP, GPipe PP).

PyTorch is the tensor/autograd layer; every hot op is a hand-written sm_100a kernel from
``csrc/`` reached through the C ABI in ``include/dtds_b200.h``; collectives are P2P / NVLS kernels
over NVLink peer memory with ``torch.distributed`` (NCCL) used for bootstrap and pipeline p2p.
"""
__version__ = "0.1.0"

PACKAGE_DIR = __path__[0]
