"""CPU suite: the C-ABI library loads and exports every symbol include/act_b200.h declares (no compute calls),
host-side logic (config validation, sharding, the missing-device failure mode) and the world_size-2 gloo path of
the data-parallel gradient averaging."""
import ctypes
import os
import re
import socket

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "act_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(act_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from act_plus_plus_b200 import _abi, _build, _lib
    _build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = _header_symbols()
    assert len(syms) >= 40
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/act_b200.h but not exported"
    # the ctypes prototype table covers the header one-to-one
    assert sorted(list(_abi.PROTOTYPES) + ["act_last_error"]) == syms
    assert _lib.lib().act_abi_version() >= 1


def test_errors_are_codes_not_exceptions():
    """No exceptions cross the ABI: a bad argument returns a negative code and act_last_error() explains it."""
    from act_plus_plus_b200 import _lib
    lib = _lib.lib()
    rc = lib.act_tc_gemm(0, 0, 0, 1, 1, None, 0, 0, 0, 0, None, 0, 0, 0, 0, None, 1, 0, 0, None)
    assert rc != 0 and len(lib.act_last_error()) > 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_path_fails_loudly_without_cuda():
    from act_plus_plus_b200._lib import ActB200Error
    from act_plus_plus_b200.engine import ActEngine
    from act_plus_plus_b200.policy import ACTPolicy
    with pytest.raises(ActB200Error):
        ActEngine(batch_max=1, img_h=64, img_w=64)
    cfg = dict(lr=1e-5, num_queries=100, kl_weight=10, hidden_dim=512, dim_feedforward=3200, lr_backbone=1e-5,
               backbone="resnet18", enc_layers=4, dec_layers=7, nheads=8, camera_names=["top"], vq=False,
               vq_class=None, vq_dim=None, action_dim=16, no_encoder=False)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ACTPolicy(cfg)


def test_unsupported_configs_raise_before_touching_the_device():
    from act_plus_plus_b200.policy import _DEFAULTS, _check_supported
    base = dict(_DEFAULTS, hidden_dim=512, nheads=8, camera_names=["top"])
    _check_supported(base)
    for bad, exc in ((dict(vq=True), NotImplementedError), (dict(use_depth=True), NotImplementedError),
                     (dict(use_pcd=True), NotImplementedError), (dict(backbone="resnet50"), ValueError),
                     (dict(pre_norm=True), ValueError), (dict(position_embedding="learned"), ValueError),
                     (dict(hidden_dim=256), ValueError)):
        with pytest.raises(exc):
            _check_supported(dict(base, **bad))


def test_shard_range_and_buckets():
    from act_plus_plus_b200.dp import shard_range, split_buckets
    for n, w in ((64, 8), (1024, 8), (10, 4), (3, 8)):
        spans = [shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1
    assert shard_range(1024, 3, 8) == (384, 512)
    b = split_buckets(100, 1000, 300)
    assert b == [(100, 300), (400, 300), (700, 300), (1000, 100)] and split_buckets(5, 0, 10) == []
    with pytest.raises(ValueError):
        shard_range(8, 8, 8)


# ------------------------------------------------------------------------------------------- gloo, world_size 2
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _dp_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from oracle import act_oracle as O
    from act_plus_plus_b200.dp import GradReducer, average_scalars, gather_env_actions, shard_range
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    cfg = O.make_cfg(num_cameras=1)
    sd = O.init_state_dict(cfg, 0)
    batch = O.make_inputs(cfg, 2, 64, 96, 0)
    lo, hi = shard_range(2, rank, world)
    shard = {k: v[lo:hi] for k, v in batch.items()}
    o = O.loss_and_grads(sd, cfg, shard)
    names = [n for n in O.trainable_names(cfg) if o["grads"][n] is not None]
    flat = torch.cat([o["grads"][n].flatten() for n in names])
    n_half = flat.numel() // 2
    red = GradReducer(flat, None, bucket_elems=3_000_000)          # several buckets per region
    red.reduce(0, n_half)                                          # "region" 1 while region 2 is still being produced
    red.reduce(n_half, flat.numel() - n_half)
    red.wait()
    losses = average_scalars(torch.stack([o["out"]["l1"], o["out"]["kl"], o["out"]["loss"]]))
    acts = gather_env_actions(torch.full((1 + rank, 3), float(rank)), [1, 2])
    if rank == 0:
        torch.save(dict(flat=flat, names=names, losses=losses, acts=acts), os.path.join(out_dir, "dp.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_owner_slices_partition_every_range():
    """Ownership map of the peer-memory optimiser step (csrc/dp_p2p.cu::dp_slice): float4 granularity, disjoint, complete."""
    from act_plus_plus_b200.dp import owner_slices
    for count in (0, 3, 4, 5, 16, 1023, 4096 + 2, 10_000_001):
        for world in (1, 2, 3, 4, 8):
            sl = owner_slices(count, world)
            cover = sorted((lo, hi) for _, lo, hi in sl)
            pos = 0
            for lo, hi in cover:
                assert lo == pos and hi > lo
                pos = hi
            assert pos == count
            sizes = [hi - lo for r, lo, hi in sl if hi <= 4 * (count // 4)]
            if sizes:
                assert max(sizes) - min(sizes) <= 4
            assert all(lo % 4 == 0 for _, lo, _ in sl)
            if count % 4:
                assert sl[-1] == (0, 4 * (count // 4), count)


def test_dp_gradient_average_equals_full_batch_gloo(tmp_path):
    """2 ranks x 1 sample, gradients averaged through GradReducer == gradients of the 2-sample batch on one rank
    (the loss is a per-rank mean over equal shards)."""
    import torch.multiprocessing as mp
    from oracle import act_oracle as O
    port = _free_port()
    mp.spawn(_dp_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = torch.load(os.path.join(tmp_path, "dp.pt"))
    cfg = O.make_cfg(num_cameras=1)
    full = O.loss_and_grads(O.init_state_dict(cfg, 0), cfg, O.make_inputs(cfg, 2, 64, 96, 0))
    ref = torch.cat([full["grads"][n].flatten() for n in got["names"]])
    rel = float((got["flat"] - ref).norm() / ref.norm())
    assert rel < 1e-4, rel
    want = torch.stack([full["out"]["l1"], full["out"]["kl"], full["out"]["loss"]])
    assert torch.allclose(got["losses"], want, rtol=1e-5)
    assert got["acts"].shape == (3, 3) and torch.equal(got["acts"][:, 0], torch.tensor([0.0, 1.0, 1.0]))


def test_batch_sampler_and_episode_pool_match_reference_contract():
    """utils.py:239-248 (BatchSampler) and :60-175 (EpisodicDataset.__getitem__): same index stream under the same numpy
    seed, and a gathered batch equals the per-item reference outputs (frames stay uint8 HWC: /255 + permute compared)."""
    import numpy as np
    import torch
    from oracle import act_oracle as O
    from act_plus_plus_b200.data import EpisodePool, batch_sampler
    rs = np.random.RandomState(0)
    K, Hh, Ww, S, A, chunk = 2, 6, 8, 14, 16, 20
    lens = [[25, 31], [18]]                                  # two dataset dirs
    episodes = []
    for i, T in enumerate([25, 31, 18]):
        episodes.append(dict(frames=rs.randint(0, 256, (T, K, Hh, Ww, 3), dtype=np.uint8), qpos=rs.randn(T, S).astype(np.float32),
                             action=rs.randn(T, A).astype(np.float32), sim=(i != 1)))
    stats = dict(qpos_mean=torch.randn(S), qpos_std=torch.rand(S) + 0.5, action_mean=torch.randn(A), action_std=torch.rand(A) + 0.5)
    np.random.seed(123)
    ref_it = O.batch_sampler_reference(5, lens, [0.7, 0.3])
    ref_batches = [next(ref_it) for _ in range(4)]
    np.random.seed(123)
    it = batch_sampler(5, lens, [0.7, 0.3])
    batches = [next(it) for _ in range(4)]
    assert ref_batches == batches
    pool = EpisodePool({k: v.numpy() for k, v in stats.items()}, chunk)
    for e in episodes:
        pool.add_episode(e["frames"], e["qpos"], e["action"], e["sim"])
    assert len(pool) == 74
    idx = batches[0] + [0, 24, 25, 55, 56, 73]              # include episode boundaries / last steps
    b = pool.gather(idx)
    assert b["image"].dtype == torch.uint8 and b["image"].shape == (len(idx), K, Hh, Ww, 3)
    for n, i in enumerate(idx):
        img, q, a, pad = O.episodic_getitem_reference(episodes, stats, chunk, i)
        assert torch.equal(b["image"][n].permute(0, 3, 1, 2).float() / 255.0, img)
        assert torch.equal(b["qpos"][n], q) and torch.equal(b["is_pad"][n], pad)
        assert torch.equal(b["actions"][n], a), (i, (b["actions"][n] - a).abs().max())
