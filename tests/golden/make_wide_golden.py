"""Mint golden vectors for hidden widths other than the reference's 64/32 (BASELINE configs[4] uses 2x256).

The reference hard-codes 64/32 (include/ppo_nn.h:85-96), so its unmodified sources cannot produce these; the
arithmetic it calls is libtorch CPU, which is the same ATen this interpreter's torch runs.  This script restates
ActorCritic::evaluate + the loss of PPO::update (src/ppo_nn.cpp:126-194, :263-290) with torch modules of the wider
shape and lets autograd produce the gradients.  tests/test_oracle_vs_golden.py holds oracle/oracle.c to them,
tests/test_gpu_wide.py holds the CUDA path to both.

usage: python tests/golden/make_wide_golden.py     (writes tests/golden/wide<H1>x<H2>_*.{f32,f64})
"""
import math
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
SIGMA, EPS_CLIP = 0.5, 0.2


def mlp(H1, H2, final_tanh):
    layers = [torch.nn.Linear(6, H1), torch.nn.Tanh(), torch.nn.Linear(H1, H2), torch.nn.Tanh(), torch.nn.Linear(H2, 1)]
    if final_tanh:
        layers.append(torch.nn.Tanh())
    return torch.nn.Sequential(*layers)


def logprob(a, mu):  # src/ppo_nn.cpp:8-54 for k = 1, same operation order
    var = torch.tensor(SIGMA * SIGMA, dtype=torch.float32)
    diff = a - mu
    numer = -0.5 * (diff * ((1.0 / var) * diff))
    denom = (torch.log(var) + math.log(2 * math.pi)) / 2
    return numer - denom


def main(H1, H2, B, seed):
    torch.manual_seed(seed)
    torch.set_num_threads(1)
    actor, critic = mlp(H1, H2, True), mlp(H1, H2, False)
    g = torch.Generator().manual_seed(seed + 1)
    th = 0.3 * torch.randn(B, generator=g)
    obs = torch.stack([torch.randn(B, generator=g), torch.randn(B, generator=g), torch.randn(B, generator=g), th,
                       torch.cos(th), torch.sin(th)], dim=1).float()
    with torch.no_grad():
        mu0 = actor(obs).squeeze(1)
        action = mu0 + SIGMA * torch.randn(B, generator=g)
        # old log-probs from a perturbed policy so that ratios leave [1-eps, 1+eps] for part of the batch
        old_logp = logprob(action, mu0 + 0.15 * torch.randn(B, generator=g))
        R = (5.0 * torch.rand(B, generator=g)).float()
    params = list(actor.parameters()) + list(critic.parameters())  # policy.parameters() order
    mu = actor(obs).squeeze(1)
    lp = logprob(action, mu)
    V = critic(obs).squeeze(1)
    ratio = torch.exp(lp - old_logp)                                  # :263
    A = R - V.detach()                                                # :274
    surr1 = ratio * A
    surr2 = torch.clamp(ratio, 1 - EPS_CLIP, 1 + EPS_CLIP) * A
    loss1 = -torch.min(surr1, surr2)                                  # :287
    loss2 = 0.5 * torch.nn.functional.mse_loss(V, R)                  # :288
    ent = 0.5 * (math.log(SIGMA * SIGMA) + (2 * math.pi * math.e) ** 1)  # :55-65 (Q8)
    loss3 = -0.01 * ent
    loss = (loss1 + loss2 + loss3).mean()                             # :290, :302
    loss.backward()
    tag = f"wide{H1}x{H2}"
    w = lambda name, t, dt: np.ascontiguousarray(t.detach().numpy().astype(dt)).tofile(os.path.join(HERE, f"{tag}_{name}"))
    w("params.f32", torch.cat([p.reshape(-1) for p in params]), np.float32)
    w("grads.f32", torch.cat([p.grad.reshape(-1) for p in params]), np.float32)
    w("obs.f32", obs, np.float32)
    w("action.f32", action, np.float32)
    w("oldlogp.f32", old_logp, np.float32)
    w("R.f32", R, np.float32)
    w("mu.f32", mu, np.float32)
    w("V.f32", V, np.float32)
    w("loss.f64", torch.tensor([loss1.mean().item(), loss2.item(), loss3, loss.item(),
                                float(((ratio < 1 - EPS_CLIP) | (ratio > 1 + EPS_CLIP)).float().mean())]), np.float64)
    print(tag, "P", sum(p.numel() for p in params), "loss", loss.item())


if __name__ == "__main__":
    main(128, 64, 384, 11)
    main(256, 256, 300, 12)
