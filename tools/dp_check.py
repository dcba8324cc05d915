"""2-rank data-parallel parity: two DP steps on shards of 8 must equal two single-process steps on the full batch of 16
(gradient averaging of equal shards == full-batch mean).  Launch with torchrun --nproc-per-node 2."""
import os
import sys
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
from act_plus_plus_b200.engine import ActEngine, R_BACKBONE, R_CVAE, R_MAIN
from act_plus_plus_b200.policy import _init_like_reference
from act_plus_plus_b200.train import train_step

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
Bs, H, W = 4, 96, 128
B = Bs * world
g = torch.Generator(device="cpu").manual_seed(7)
full = dict(image=torch.rand((B, 1, 3, H, W), generator=g), qpos=torch.randn((B, 14), generator=g),
            actions=torch.randn((B, 100, 16), generator=g), is_pad=(torch.arange(100)[None].repeat(B, 1) >= 70),
            eps=torch.randn((B, 32), generator=g))
full = {k: v.cuda() for k, v in full.items()}
mine = {k: v[rank * Bs:(rank + 1) * Bs].contiguous() for k, v in full.items()}
eng = ActEngine(batch_max=Bs, img_h=H, img_w=W, train=True)
_init_like_reference(eng, seed=3)
eng.set_lr(1e-3, 1e-3)
P2P = os.environ.get("ACT_DP_P2P", "0") == "1"
if P2P:
    eng.dp_connect(dist.group.WORLD)       # gradient exchange inside the optimiser kernel over peer memory (csrc/dp_p2p.cu)
    if rank == 0:
        print("dp_check exchange:", "NVLS multicast" if eng._dp_nvls else f"peer loads/stores ({getattr(eng, '_dp_nvls_error', 'multicast off')})", flush=True)
NSTEPS = int(os.environ.get('ACT_DP_STEPS', '2'))
snaps = []
for _ in range(NSTEPS):
    train_step(eng, mine["qpos"], mine["image"], mine["actions"], mine["is_pad"], mine["eps"], dist.group.WORLD)
    snaps.append(eng.flat_params.clone())
torch.cuda.synchronize()
ok = True
if P2P:
    ok = eng.dp_status() == 0
    print(f"dp_check p2p rank {rank}: barrier status {'OK' if ok else 'TIMEOUT'}, {len(eng.dp_slices())} ranges exchanged")
if rank == 0:
    ref = ActEngine(batch_max=B, img_h=H, img_w=W, train=True)
    _init_like_reference(ref, seed=3)
    p0 = ref.flat_params.clone()
    ref.set_lr(1e-3, 1e-3)
    for i in range(NSTEPS):
        train_step(ref, full["qpos"], full["image"], full["actions"], full["is_pad"], full["eps"])
        torch.cuda.synchronize()
        for name, r in (("main", R_MAIN), ("cvae", R_CVAE), ("backbone", R_BACKBONE)):
            off, n = ref.region(r)
            d = (snaps[i][off:off + n] - ref.flat_params[off:off + n]).abs()
            moved = (ref.flat_params[off:off + n] - p0[off:off + n]).abs().mean().item()
            frac = (d > 0.5 * moved).float().mean().item()
            # After step 1 every region must agree with the full-batch run (first Adam step = lr * sign(g): only elements
            # whose gradient is within bf16 noise of zero may differ).  At step 2 the backbone gradient of this tiny
            # random-init problem has shrunk 40x (|g| 1.6 -> 0.04) and is noise-dominated in BOTH runs (measured: its
            # cosine to the full-batch gradient varies 0.35-0.9 between identical launches of either exchange path), so
            # only a loose bound is meaningful there.
            loose = name == "backbone" and i >= 1
            good = moved > 2e-4 and d.mean().item() <= (0.15 if loose else 0.05) * moved and frac < (1e-1 if loose else 2e-2)
            ok &= good
            print(f"dp_check step {i + 1} {name}: moved {moved:.3e} mean|dp-full| {d.mean().item():.3e} diverging {frac:.2e} {'OK' if good else 'FAIL'}")
if os.environ.get("ACT_DP_DIAG", "0") == "1":
    # p2p: local gradients survive the step, average them here; NCCL: the buffer already holds the average.  Compare with
    # the full-batch gradient per region.
    if P2P:
        gl = [torch.zeros_like(eng.flat_grads) for _ in range(world)]
        dist.all_gather(gl, eng.flat_grads)
        gavg = sum(gl) / world
    else:
        gavg = eng.flat_grads
    if rank == 0:
        for nm, r in (("main", R_MAIN), ("cvae", R_CVAE), ("backbone", R_BACKBONE)):
            o, n = ref.region(r)
            ga, gr = gavg[o:o + n], ref.flat_grads[o:o + n]
            print("dp_check diag %-9s grad relerr %.3e |g_dp| %.3e |g_full| %.3e cos %.4f" % (
                nm, ((ga - gr).norm() / (gr.norm() + 1e-30)).item(), ga.norm().item(), gr.norm().item(),
                (torch.dot(ga, gr) / (ga.norm() * gr.norm() + 1e-30)).item()))
t = torch.tensor([1.0 if ok else 0.0], device="cuda")
dist.broadcast(t, 0)
# all ranks must hold identical parameters after the averaged update
chk = eng.flat_params.double().sum().reshape(1)
lst = [torch.zeros_like(chk) for _ in range(world)]
dist.all_gather(lst, chk)
same = all(torch.equal(lst[0], x) for x in lst)
if P2P:
    # the bf16 shadows (tensor-core operands) must be identical too: compare a second forward's losses across ranks on
    # the SAME inputs (rank 0's shard)
    shard0 = {k: v[:Bs].contiguous() for k, v in full.items()}
    l = eng.forward(shard0["qpos"], shard0["image"], shard0["actions"], shard0["is_pad"], shard0["eps"])["losses"].double().reshape(1, 3)
    ll = [torch.zeros_like(l) for _ in range(world)]
    dist.all_gather(ll, l)
    same_fwd = all(torch.equal(ll[0], x) for x in ll)
    same = same and same_fwd
    # the derived operands (bf16 shadows written by the peers, conv packs rebuilt after the broadcast) must equal what a
    # fresh derivation from the fp32 masters gives
    eng.refresh()
    l2 = eng.forward(shard0["qpos"], shard0["image"], shard0["actions"], shard0["is_pad"], shard0["eps"])["losses"].double().reshape(1, 3)
    fresh = torch.equal(l, l2)
    same = same and fresh
    print(f"dp_check p2p rank {rank}: derived operands consistent with fp32 masters: {fresh} {l.tolist()} {l2.tolist()}")
    if rank == 0:
        print("dp_check p2p forward identical across ranks:", same_fwd)
    # sharded moments -> complete on every rank
    eng.dp_gather_moments(dist.group.WORLD)
    mchk = eng.exp_avg.double().abs().sum().reshape(1)
    ml = [torch.zeros_like(mchk) for _ in range(world)]
    dist.all_gather(ml, mchk)
    same = same and all(torch.equal(ml[0], x) for x in ml)
if rank == 0:
    print("dp_check replicas identical:", same)
torch.cuda.synchronize()
sys.stdout.flush()
os._exit(0 if (t.item() == 1.0 and same) else 1)
