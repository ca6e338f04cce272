import sys, copy, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from test_har_native import _setup, _rel
for B in (32, 128):
    mod, lay, row, tr, x, y = _setup(B=B, seed=9)
    h = mod.conv(x[:, None, :]).transpose(1, 2)
    h = mod.pe(h).transpose(0, 1)
    enc = mod.transformer(h); enc.retain_grad()
    pooled = enc.mean(dim=0); pooled.retain_grad()
    c1pre = mod.classifier[0](pooled); c1pre.retain_grad()
    logits = mod.classifier[3](mod.classifier[2](mod.classifier[1](c1pre)))
    loss = torch.nn.functional.cross_entropy(logits, y); loss.backward()
    grad = torch.zeros_like(row); tr.out3.zero_()
    tr.grad_step(row, grad, x, y, dropout=False)
    got = lay.views(grad)
    M = B * 561
    print(B, "pooled", _rel(tr.pooled[:B], pooled), "logits", _rel(tr.logits[:B], logits), "dpooled", _rel(tr.dpooled[:B], pooled.grad))
    w1 = mod.classifier[0].weight; b1 = mod.classifier[0].bias
    pre_n = tr.pooled[:B] @ w1.t() + b1
    flips = ((pre_n > 0) != (c1pre > 0)).sum().item()
    print("   relu flips in head:", flips, "min |pre|", c1pre.abs().min().item())
    # recompute head grads in torch from the native pooled to separate head-kernel errors from mask flips
    pn = tr.pooled[:B].detach().clone().requires_grad_(True)
    lg = mod.classifier[3](torch.relu(mod.classifier[0](pn)))
    mod.zero_grad(); torch.nn.functional.cross_entropy(lg, y).backward()
    print("   classifier.0.weight vs torch-on-native-pooled", _rel(got["classifier.0.weight"], mod.classifier[0].weight.grad),
          "dpooled", _rel(tr.dpooled[:B], pn.grad))
    enc_n = tr.layers[1]["h2"][:M].reshape(B, 561, 64).transpose(0, 1)
    print("   encoder out", _rel(enc_n, enc))
