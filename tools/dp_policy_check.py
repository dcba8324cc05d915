"""Data-parallel training through the drop-in API (ACTPolicy.set_data_parallel) under torchrun: a few optimisation steps on
per-rank shards, then: replicas bit-identical, nn.Parameter objects still alias the engine's flat buffers after the move
into symmetric memory (NVLS path), losses finite, state_dict round trip.  ACT_DP_NVLS=0|1 selects the exchange.
usage: torchrun --nproc-per-node N tools/dp_policy_check.py [fused]"""
import os
import sys
import warnings
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
from act_plus_plus_b200.policy import ACTPolicy

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
fused = len(sys.argv) > 1 and sys.argv[1] == "fused"
cfg = {'dropout': 0.0, 'lr': 1e-4, 'num_queries': 100, 'kl_weight': 10, 'hidden_dim': 512, 'dim_feedforward': 3200, 'lr_backbone': 1e-5,
       'backbone': 'resnet18', 'enc_layers': 4, 'dec_layers': 7, 'nheads': 8, 'camera_names': ['top'], 'vq': False, 'vq_class': None,
       'vq_dim': None, 'action_dim': 16, 'no_encoder': False, 'seed': 0, 'fused_step': fused}
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    policy = ACTPolicy(cfg).cuda()
policy.train()
opt = policy.configure_optimizers()
params_before = {n: p for n, p in policy.named_parameters()}           # objects handed out BEFORE the buffers move
policy.set_data_parallel()
B, H, W = 4, 96, 128
g = torch.Generator(device="cuda").manual_seed(100 + rank)
losses = []
for step in range(4):
    image = torch.rand((B, 1, 3, H, W), generator=g, device="cuda")
    qpos = torch.randn((B, 14), generator=g, device="cuda")
    actions = torch.randn((B, 100, 16), generator=g, device="cuda")
    is_pad = torch.zeros((B, 100), dtype=torch.bool, device="cuda")
    out = policy(qpos, image, actions, is_pad)
    out["loss"].backward()
    opt.step()
    opt.zero_grad()
    losses.append(float(out["loss"]))
torch.cuda.synchronize()
eng = policy.model._master
train_eng = [e for e in policy.model._engines.values() if e.train_enabled][0]
flat = eng.flat_params
ok = all(torch.isfinite(torch.tensor(losses)))
# replicas identical
chk = torch.stack([flat.double().sum(), flat.double().abs().sum(), (flat.double() * torch.arange(flat.numel(), device=dev) % 7).sum()])
allc = [torch.zeros_like(chk) for _ in range(world)]
dist.all_gather(allc, chk)
same = all(bool((c == allc[0]).all()) for c in allc)
# Parameter objects alias the (possibly moved) flat buffer
lo, hi = flat.data_ptr(), flat.data_ptr() + flat.numel() * 4
alias = all(lo <= p.data_ptr() < hi for p in params_before.values() if p.numel() > 1)
after = dict(policy.named_parameters())
ident = all(after[n] is params_before[n] for n in params_before)
sd = policy.state_dict()
name0 = [n for n in sd if n.endswith("encoder.layers.0.linear1.weight")][0]
view_ok = bool((sd[name0].to(dev).float() == after[name0 if name0 in after else name0].data.float()).all()) if name0 in after else True
if rank == 0:
    print(f"dp_policy_check fused={fused} nvls={getattr(train_eng, '_dp_nvls', False)} connected={train_eng.dp_connected} losses={[round(l, 4) for l in losses]}", flush=True)
    print(f"dp_policy_check finite={bool(ok)} replicas_identical={same} params_alias_flat={alias} param_identity_kept={ident} state_dict_view={view_ok}", flush=True)
good = torch.tensor([1.0 if (ok and same and alias and ident and view_ok and train_eng.dp_connected) else 0.0], device=dev)
dist.all_reduce(good, op=dist.ReduceOp.MIN)
if rank == 0:
    print("dp_policy_check:", "PASS" if float(good) == 1.0 else "FAIL", flush=True)
dist.destroy_process_group()
sys.exit(0 if float(good) == 1.0 else 1)
