import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dtds_b200 import comm
from dtds_b200.models import build_model, get_config
from dtds_b200.parallel import zero
rank = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(rank)
comm.init_process_group()
arch = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("--") else "bert-tiny"
stages = [int(a) for a in sys.argv[2:] if a.isdigit()] or [0, 1, 2, 3]
cfg = get_config(arch).replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
ids = torch.randint(0, cfg.vocab_size, (2, 128), generator=torch.Generator().manual_seed(11 + rank))
for stage in stages:
    torch.manual_seed(0)
    model = build_model(cfg)
    conf = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
            "zero_optimization": {"stage": stage, "reduce_bucket_size": 2e5}, "bf16": {"enabled": True}}
    eng, opt, _, _ = zero.initialize(model=model, model_parameters=model.parameters(), config=conf)
    for it in range(3):
        loss = eng(ids.cuda(), labels=ids.cuda()).loss
        torch.cuda.synchronize(); print(rank, "stage", stage, "it", it, "fwd ok", loss.item(), flush=True)
        eng.backward(loss)
        torch.cuda.synchronize(); print(rank, "stage", stage, "it", it, "bwd ok", flush=True)
        eng.step()
        torch.cuda.synchronize(); print(rank, "stage", stage, "it", it, "step ok", flush=True)
comm.destroy_process_group()
