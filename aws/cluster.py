"""Cluster description shared by the CDK stacks and the shell scripts: pure Python, no AWS imports,
so the hostfile / ssh-config rendering is unit-tested on any machine.

The reference provisions T4 boxes (one g4dn.12xlarge, or three g4dn.xlarge with one GPU each:
aws/training_stack.py:20, aws/multi_node_training_stack.py:17-19).  The unit here is one
8×B200 NVSwitch node (p6-b200.48xlarge): the peer-memory / NVLS kernels work inside such a node,
several nodes talk NCCL over EFA.
"""
from __future__ import annotations

from dataclasses import dataclass

B200_INSTANCE_TYPE = "p6-b200.48xlarge"
GPUS_PER_NODE = 8
AMI_NAME_PATTERN = "Deep Learning Base OSS Nvidia Driver GPU AMI (Ubuntu 22.04)*"
LOGIN_USER = "ubuntu"


@dataclass(frozen=True)
class ClusterSpec:
    nodes: int = 1
    instance_type: str = B200_INSTANCE_TYPE
    gpus_per_node: int = GPUS_PER_NODE
    ami_name: str = AMI_NAME_PATTERN
    user: str = LOGIN_USER

    def worker_names(self) -> list[str]:
        return [f"worker-{i + 1}" for i in range(self.nodes)]


def render_hostfile(dns_names: list[str], slots: int = GPUS_PER_NODE) -> str:
    """DeepSpeed-syntax hostfile, one ``<host> slots=<gpus>`` line per node — what
    ``python -m dtds_b200.launch --hostfile ... --ssh`` reads."""
    return "".join(f"{name} slots={slots}\n" for name in dns_names)


def render_ssh_config(dns_names: list[str], user: str = LOGIN_USER, key_file: str = "{YOUR_SSH_KEY_FILE_PATH}") -> str:
    """``Host worker-N`` blocks for the operator's ~/.ssh/config."""
    blocks = []
    for i, name in enumerate(dns_names):
        blocks.append(f"Host worker-{i + 1}\n    HostName {name}\n    User {user}\n"
                      f"    IdentityFile {key_file}\n    StrictHostKeyChecking no\n")
    return "\n".join(blocks)


def connect_command(key_name: str, dns_name: str, user: str = LOGIN_USER) -> str:
    return f"ssh -i {key_name}.pem {user}@{dns_name}"
