import copy, torch, sys
sys.path.insert(0, "/root/repo")
from dtds_b200.models import build_model, get_config
name = sys.argv[1] if len(sys.argv) > 1 else "bloom-tiny"
torch.manual_seed(0)
cfg = get_config(name).replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
ref = build_model(cfg).float()
gpu = copy.deepcopy(ref).to(device="cuda", dtype=torch.bfloat16)
with torch.no_grad():
    for pr, pg in zip(ref.parameters(), gpu.parameters()):
        pr.copy_(pg.float().cpu())
ids = torch.randint(0, cfg.vocab_size, (2, 128)); labels = ids.clone()
lr = ref(ids, labels=labels).loss; lr.backward()
lg = gpu(ids.cuda(), labels=labels.cuda()).loss; lg.backward()
print("loss", lr.item(), lg.item())
for (n, pr), pg in zip(ref.named_parameters(), gpu.parameters()):
    gr, gg = pr.grad.float(), pg.grad.float().cpu()
    print(f"{n:40s} refmax {gr.abs().max():.3e} relerr {(gr-gg).abs().max()/(gr.abs().max()+1e-9):.3f} cos {torch.nn.functional.cosine_similarity(gr.flatten(), gg.flatten(), dim=0):.4f}")
