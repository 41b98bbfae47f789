"""torch-CPU autograd transcription of the reference loss/update graph.

Used to self-check the NumPy oracle (tests only).  Mirrors dqn/agent.py:85-117,
double_q/agent.py:85-111 and dqn_mgsc_batched/agent.py:104-220 literally, with
autograd doing the differentiation the reference delegates to jax.grad.
"""
import numpy as np
import torch
import torch.nn.functional as F


class ClipGradient(torch.autograd.Function):
    """rlax.clip_gradient: identity forward, clip the cotangent backward."""

    @staticmethod
    def forward(ctx, x, lo, hi):
        ctx.lo, ctx.hi = lo, hi
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g.clamp(ctx.lo, ctx.hi), None, None


def to_torch(params, dtype=torch.float64, requires_grad=False):
    return {k: torch.tensor(np.asarray(v), dtype=dtype, requires_grad=requires_grad)
            for k, v in params.items()}


def forward(p, s_u8):
    dt = p["fc1.w"].dtype
    x = torch.as_tensor(s_u8).to(dt) / 255.0           # NHWC
    x = x.permute(0, 3, 1, 2)                           # NCHW for torch
    for name, s in (("conv1", 4), ("conv2", 2), ("conv3", 1)):
        w = p[name + ".w"].permute(3, 2, 0, 1)          # HWIO -> OIHW
        x = F.relu(F.conv2d(x, w, p[name + ".b"], stride=s))
    h0 = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)  # flatten in (h,w,c) order
    h1 = F.relu(h0 @ p["fc1.w"] + p["fc1.b"])
    return h1 @ p["fc2.w"] + p["fc2.b"]


def loss_fn(online, target, batch, double_q=False, grad_bound=1.0 / 32, weights=None):
    s_tm1, a, r, disc, s_t = batch
    dt = online["fc1.w"].dtype
    a = torch.as_tensor(np.asarray(a), dtype=torch.long)
    r = torch.as_tensor(np.asarray(r), dtype=dt)
    disc = torch.as_tensor(np.asarray(disc), dtype=dt)
    q_tm1 = forward(online, s_tm1)
    with torch.no_grad():
        q_tgt = forward(target, s_t)
        sel = (forward(online, s_t) if double_q else q_tgt).argmax(1)
        y = r + disc * q_tgt.gather(1, sel[:, None])[:, 0]
    td = y - q_tm1.gather(1, a[:, None])[:, 0]
    td = ClipGradient.apply(td, -grad_bound, grad_bound)
    losses = 0.5 * td * td
    if weights is not None:     # prioritized/agent.py:103-108: jnp.mean(losses * weights)
        losses = losses * torch.as_tensor(np.asarray(weights), dtype=dt)
    return losses.mean(), q_tm1


def grads(online_np, target_np, batch, double_q=False, grad_bound=1.0 / 32, dtype=torch.float64, weights=None):
    online = to_torch(online_np, dtype, requires_grad=True)
    target = to_torch(target_np, dtype)
    loss, q = loss_fn(online, target, batch, double_q, grad_bound, weights)
    g = torch.autograd.grad(loss, list(online.values()))
    return float(loss), q.detach().numpy(), {k: gi.numpy() for k, gi in zip(online, g)}


def _flat(d):
    return torch.cat([v.reshape(-1) for v in d.values()])


def _unflat(flat, like):
    out, o = {}, 0
    for k, v in like.items():
        n = v.numel()
        out[k] = flat[o:o + n].reshape(v.shape)
        o += n
    return out


def _rmsprop(theta, g, mu, nu, lr, decay, eps):
    mu2 = (1 - decay) * g + decay * mu
    nu2 = (1 - decay) * g * g + decay * nu
    return theta + (-lr) * g * torch.rsqrt(nu2 - mu2 * mu2 + eps), mu2, nu2


def hvp(online_np, target_np, batch, direction_np, double_q=False, grad_bound=1.0 / 32, weights=None):
    """Hessian-vector product by double backward through loss_fn (clip_gradient's backward is differentiable)."""
    online = to_torch(online_np, torch.float64, requires_grad=True)
    target = to_torch(target_np, torch.float64)
    loss, _ = loss_fn(online, target, batch, double_q, grad_bound, weights)
    g = torch.autograd.grad(loss, list(online.values()), create_graph=True)
    d = to_torch(direction_np, torch.float64)
    inner = sum((gi * d[k]).sum() for gi, k in zip(g, online))
    hv = torch.autograd.grad(inner, list(online.values()))
    return {k: x.numpy() for k, x in zip(online, hv)}


def meta_grad(logits_np, meta_batch, online_np, target_np, online_transition, mu_np, nu_np,
              grad_bound=1.0 / 32, lr=0.00025, decay=0.95, eps=0.01 / 32 ** 2, second_order=False):
    """d meta_loss / d logits by autograd through a literal transcription of meta_loss_fn."""
    dt = torch.float64
    online = to_torch(online_np, dt, requires_grad=True)
    target = to_torch(target_np, dt)
    logits = torch.tensor(np.asarray(logits_np), dtype=dt, requires_grad=True)
    mu, nu = torch.tensor(mu_np, dtype=dt), torch.tensor(nu_np, dtype=dt)
    s_tm1, a, r, disc, s_t = meta_batch
    c = logits.max()
    p = torch.exp(logits - (c + torch.log(torch.sum(torch.exp(logits - c)))))
    weighted = []
    for i in range(len(a)):
        one = (s_tm1[i:i + 1], a[i:i + 1], r[i:i + 1], disc[i:i + 1], s_t[i:i + 1])
        li, _ = loss_fn(online, target, one, False, grad_bound)
        gi = _flat(dict(zip(online, torch.autograd.grad(li, list(online.values())))))
        weighted.append(p[i] * gi.detach())
    G = torch.stack(weighted).sum(0)
    theta = _flat(online).detach()
    theta_exp, mu1, nu1 = _rmsprop(theta, G, mu, nu, lr, decay, eps)
    if second_order:      # dqn_mgsc_batched_reservoir/agent.py:183-195: no stop_gradient anywhere
        one = tuple(np.asarray(x)[None, ...] for x in online_transition)
        l2, _ = loss_fn(_unflat(theta_exp, online), {k: v.detach() for k, v in online.items()}, one, False, grad_bound)
        g2 = torch.autograd.grad(l2, theta_exp, create_graph=True)[0]
        theta_tgt, _, _ = _rmsprop(theta_exp, g2, mu1, nu1, lr, decay, eps)
        L = torch.sum((theta_exp - theta_tgt) ** 2)
        (dlogits,) = torch.autograd.grad(L, logits)
        return float(L), dlogits.numpy()
    # gradient at the expected params with the *online* params as target net (agent.py:183-185)
    exp_leaf = theta_exp.detach().clone().requires_grad_(True)
    one = tuple(np.asarray(x)[None, ...] for x in online_transition)
    l2, _ = loss_fn(_unflat(exp_leaf, online), {k: v.detach() for k, v in online.items()}, one,
                    False, grad_bound)
    g2 = torch.autograd.grad(l2, exp_leaf)[0]
    theta_tgt, _, _ = _rmsprop(theta_exp.detach(), g2, mu1.detach(), nu1.detach(), lr, decay, eps)
    L = torch.sum((theta_exp - theta_tgt.detach()) ** 2)
    (dlogits,) = torch.autograd.grad(L, logits)
    return float(L), dlogits.numpy()
