#!/usr/bin/env python3
"""CDK app: `TrainingStack` (one 8×B200 node) and `MultiNodeTrainingStack` (TRAINING_NODES nodes,
default 2).  Account / region / key pair come from the environment, like the reference's app
(aws/app.py:14-17): CDK_DEFAULT_ACCOUNT, CDK_DEFAULT_REGION, EC2_SSH_KEY_NAME.

    cd aws && cdk deploy TrainingStack
"""
import os

import aws_cdk as cdk

from cluster import ClusterSpec
from training_stack import build_training_stack

app = cdk.App()
env = cdk.Environment(account=os.getenv("CDK_DEFAULT_ACCOUNT"), region=os.getenv("CDK_DEFAULT_REGION"))
key = os.getenv("EC2_SSH_KEY_NAME")
build_training_stack(app, "TrainingStack", env, key, ClusterSpec(nodes=1))
build_training_stack(app, "MultiNodeTrainingStack", env, key, ClusterSpec(nodes=int(os.getenv("TRAINING_NODES", "2"))))
app.synth()
