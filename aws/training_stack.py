"""One-node and N-node training stacks for 8×B200 instances (AWS CDK v2).

Counterpart of the reference's two stacks (aws/training_stack.py, aws/multi_node_training_stack.py)
with one class: ``nodes=1`` is the single box every benchmark of this repository runs on;
``nodes>1`` adds what several nodes need — a cluster placement group, a security group that allows
all traffic between the nodes (NCCL bootstrap and EFA), and the hostfile / ssh-config outputs that
``scripts/setup-workers.sh`` and ``scripts/launch-multinode.sh`` consume.

aws-cdk-lib is imported lazily: the module can be imported (and `aws/cluster.py` tested) without it.
"""
from __future__ import annotations

from cluster import ClusterSpec, connect_command, render_hostfile, render_ssh_config


def build_training_stack(scope, construct_id: str, env, ssh_key_name: str, spec: ClusterSpec | None = None):
    from aws_cdk import CfnOutput, Stack, aws_ec2 as ec2

    spec = spec or ClusterSpec()
    stack = Stack(scope, construct_id, env=env)
    vpc = ec2.Vpc(stack, "TrainingVpc", max_azs=1, subnet_configuration=[
        ec2.SubnetConfiguration(name="TrainingSubnet", subnet_type=ec2.SubnetType.PUBLIC)])
    group = ec2.SecurityGroup(stack, "TrainingNodes", vpc=vpc, allow_all_outbound=True)
    group.add_ingress_rule(ec2.Peer.any_ipv4(), ec2.Port.tcp(22), "ssh")
    if spec.nodes > 1:
        group.add_ingress_rule(group, ec2.Port.all_traffic(), "node to node (NCCL bootstrap, EFA)")
    placement = ec2.PlacementGroup(stack, "TrainingPlacement",
                                   strategy=ec2.PlacementGroupStrategy.CLUSTER) if spec.nodes > 1 else None
    image = ec2.MachineImage.lookup(name=spec.ami_name)
    names = []
    for i in range(spec.nodes):
        node = ec2.Instance(stack, f"TrainingNode{i + 1}", vpc=vpc, instance_type=ec2.InstanceType(spec.instance_type),
                            machine_image=image, security_group=group, key_name=ssh_key_name,
                            placement_group=placement,
                            block_devices=[ec2.BlockDevice(device_name="/dev/sda1",
                                                           volume=ec2.BlockDeviceVolume.ebs(500))])
        names.append(node.instance_public_dns_name)
    if spec.nodes == 1:
        CfnOutput(stack, "ConnectCommand", value=connect_command(ssh_key_name, names[0], spec.user))
    else:
        CfnOutput(stack, "HostfileContent", value="\n" + render_hostfile(names, spec.gpus_per_node))
        CfnOutput(stack, "SshConfigContent", value="\n" + render_ssh_config(names, spec.user))
    return stack
