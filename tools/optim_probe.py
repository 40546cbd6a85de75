import copy, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import llm_fp8_b200 as te
from llm_fp8_b200.layer import LinearPathStack
te.reset_fp8_state()
model = LinearPathStack("llama-3.2-1b", "mxfp8", 1, device="cuda", seed=5)
twin = copy.deepcopy(model)
init = {n: p.detach().clone() for n, p in model.named_parameters()}
opt = te.FusedAdamW(model, lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1)
ref = torch.optim.AdamW(twin.parameters(), lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1)
print("fused entries:", [(tuple(w.shape), len(ps), key) for w, ps, _, key in opt.fused])
g = torch.Generator().manual_seed(1)
for step in range(3):
    x = torch.randn(2, 128, model.H, generator=g).bfloat16().cuda()
    grads = {}
    for name, net, o in (("fused", model, opt), ("torch", twin, ref)):
        o.zero_grad()
        net(x).float().pow(2).mean().backward()
        grads[name] = {n: p.grad.detach().float().clone() for n, p in net.named_parameters()}
        o.step()
    for (n, p), (_, q) in zip(model.named_parameters(), twin.named_parameters()):
        if p.dim() < 2:
            continue
        gd = (grads["fused"][n] - grads["torch"][n]).abs().max().item() / (grads["torch"][n].abs().max().item() + 1e-30)
        df, dt = (p.float() - init[n].float()), (q.float() - init[n].float())
        print(f"step {step} {n:34s} |g| {grads['torch'][n].abs().mean().item():.2e} grad rel diff {gd:.2e}  mean|dw| fused {df.abs().mean().item():.2e} torch {dt.abs().mean().item():.2e}  max|w_f - w_t| {(p.float()-q.float()).abs().max().item():.2e}  sign disagree {((df.sign() != dt.sign()) & (dt != 0)).float().mean().item():.3f}")
