"""One training step of BASELINE.json configs[2]: torchvision ResNet-18 + dnn_to_bnn(type="Flipout"), synthetic
3x224x224, batch 128: forward, get_kl_loss, CE + kl/batch, backward, SGD step
(the loop of examples/main_bayesian_cifar_dnn2bnn.py:374-449 with num_mc = 1).  bf16 channels-last activations,
fp32 variational parameters.  Prints forward-only and whole-step times.
usage: train_step.py [batch] [iters] [type]"""
import os
import sys

import torch
import torchvision

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
from bayesian_torch_b200 import ops  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 128
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
typ = sys.argv[3] if len(sys.argv) > 3 else "Flipout"
arch = os.environ.get("BT_ARCH", "resnet18")
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = getattr(torchvision.models, arch)()
bt.dnn_to_bnn(model, {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
                      "type": typ, "moped_enable": False, "moped_delta": 0.5})
model = model.to(dev).bfloat16().to(memory_format=torch.channels_last).train()
opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
x = torch.randn(batch, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 1000, (batch,), device=dev)


def step():
    opt.zero_grad(set_to_none=True)
    out = model(x)
    kl = bt.get_kl_loss(model)
    loss = torch.nn.functional.cross_entropy(out.float(), y) + kl / batch
    loss.backward()
    opt.step()
    return loss


def timed(fn, n):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        r = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, r


with torch.no_grad():
    fwd_ms, _ = timed(lambda: model(x), iters)
ops.PROFILE.reset()
step_ms, loss = timed(step, iters)
print(f"{arch} dnn_to_bnn({typ}) batch {batch}: forward {fwd_ms:.2f} ms ({batch / fwd_ms * 1e3:.0f} images/s), "
      f"training step {step_ms:.2f} ms ({batch / step_ms * 1e3:.0f} images/s), loss {float(loss):.4f}, "
      f"library launches per step {ops.PROFILE.launches // (iters + 1)}")
