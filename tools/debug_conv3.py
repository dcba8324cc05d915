import sys, torch
sys.path.insert(0, ".")
from act_plus_plus_b200 import ops
torch.manual_seed(0)
N, H, W, Cin, Cout = 1, 12, 40, 64, 64
x = torch.randn(N, H, W, Cin, device="cuda").bfloat16()
for tap in list(range(9)) + [-1]:
    w = torch.zeros(Cout, Cin, 3, 3, device="cuda")
    if tap >= 0:
        w[:, :, tap // 3, tap % 3] = torch.randn(Cout, Cin, device="cuda") * 0.1
    else:
        w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.1
    wpk, _ = ops.pack_conv_weights(w)
    y = ops.conv_fwd(x, wpk, Cout, 3, 1, 1)
    ref = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.bfloat16().float(), padding=1).permute(0, 2, 3, 1)
    err = (y.float() - ref).abs()
    rel = float(err.norm() / ref.norm())
    bad = (err > 0.05).any(dim=-1)[0]          # [H, W] mask of wrong pixels
    rows = bad.any(dim=1).nonzero().flatten().tolist()
    cols = bad.any(dim=0).nonzero().flatten().tolist()
    print(f"tap {tap}: rel {rel:.4f} bad pixels {int(bad.sum())}/{H*W} rows {rows[:12]} cols {cols[:16]}")
