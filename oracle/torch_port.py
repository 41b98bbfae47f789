"""TEST / BASELINE INFRASTRUCTURE ONLY — torch-CPU fp32 port of the reference `update`.

The reference's learner runs on JAX (`--jax_platform_name=cpu` for its CPU path); JAX is
not installable here, so the timed CPU baseline (bench.py `cpu_baseline`, `--impl reference`)
is this port: the same graph (dqn/agent.py:85-117, double_q/agent.py:85-111,
optax.rmsprop centered) on torch's oneDNN CPU kernels with all host threads.  It is checked
against the float64 oracle in tests/test_torch_port.py.  kind = "port".
"""
import numpy as np
import torch
import torch.nn.functional as F

from oracle import learner_oracle as net


class _ClipGrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bound):
        ctx.bound = bound
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g.clamp(-ctx.bound, ctx.bound), None


class TorchCpuLearner:
    def __init__(self, params_np, double_q=False, grad_bound=1.0 / 32, lr=0.00025, decay=0.95,
                 eps=0.01 / 32 ** 2, threads=None):
        if threads:
            torch.set_num_threads(threads)
        self.threads = torch.get_num_threads()
        self.double_q, self.bound, self.lr, self.decay, self.eps = double_q, grad_bound, lr, decay, eps
        # torch-native layouts (OIHW conv weights) so oneDNN sees what it likes
        self.online = self._to_torch(params_np, True)
        self.target = self._to_torch(params_np, False)
        self.mu = [torch.zeros_like(p) for p in self.online.values()]
        self.nu = [torch.zeros_like(p) for p in self.online.values()]

    @staticmethod
    def _to_torch(params_np, grad):
        out = {}
        for k, v in params_np.items():
            t = torch.tensor(np.asarray(v), dtype=torch.float32)
            if k.startswith("conv") and k.endswith(".w"):
                t = t.permute(3, 2, 0, 1).contiguous()
            out[k] = t.requires_grad_(grad)
        return out

    def sync_target(self):
        with torch.no_grad():
            for k in self.target:
                self.target[k].copy_(self.online[k])

    @staticmethod
    def _forward(p, s_u8):
        x = torch.from_numpy(s_u8).permute(0, 3, 1, 2).to(torch.float32) / 255.0
        for name, s in (("conv1", 4), ("conv2", 2), ("conv3", 1)):
            x = F.relu(F.conv2d(x, p[name + ".w"], p[name + ".b"], stride=s))
        h0 = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        h1 = F.relu(h0 @ p["fc1.w"] + p["fc1.b"])
        return h1 @ p["fc2.w"] + p["fc2.b"]

    def update(self, s_tm1, a, r, disc, s_t):
        a = torch.as_tensor(np.asarray(a), dtype=torch.long)
        r = torch.as_tensor(np.asarray(r), dtype=torch.float32)
        disc = torch.as_tensor(np.asarray(disc), dtype=torch.float32)
        q_tm1 = self._forward(self.online, s_tm1)
        with torch.no_grad():
            q_tgt = self._forward(self.target, s_t)
            sel = (self._forward(self.online, s_t) if self.double_q else q_tgt).argmax(1)
            y = r + disc * q_tgt.gather(1, sel[:, None])[:, 0]
        td = _ClipGrad.apply(y - q_tm1.gather(1, a[:, None])[:, 0], self.bound)
        loss = (0.5 * td * td).mean()
        grads = torch.autograd.grad(loss, list(self.online.values()))
        with torch.no_grad():
            for p, g, m, v in zip(self.online.values(), grads, self.mu, self.nu):
                m.mul_(self.decay).add_(g, alpha=1 - self.decay)
                v.mul_(self.decay).addcmul_(g, g, value=1 - self.decay)
                p.add_(g * torch.rsqrt(v - m * m + self.eps), alpha=-self.lr)
        return float(loss)

    def params_np(self):
        out = {}
        for k, t in self.online.items():
            t = t.detach()
            if k.startswith("conv") and k.endswith(".w"):
                t = t.permute(2, 3, 1, 0)
            out[k] = t.numpy().copy()
        return out


def make(num_actions, shared_bias, double_q, seed=0, threads=None):
    params = net.init_params(np.random.default_rng(seed), num_actions, shared_bias)
    return TorchCpuLearner(params, double_q=double_q, threads=threads)
