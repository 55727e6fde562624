"""ORACLE -- TEST INFRASTRUCTURE ONLY.  torch-CPU restatement of the step with autograd.

Two uses: (1) in fp64, torch.autograd judges the hand-written backward of oracle/step_np.py
(tests/test_oracle.py); (2) in fp32 with all host threads it is the timed CPU baseline that stands
in for the reference's CPU TensorFlow path, which is unavailable on this box:
``bench.py`` cpu_baseline and ``bench.py --impl reference``.  kind = "port".

Follows the same reference lines as step_np.py: masked.py:110-160 (closed form: dilated F.conv1d with
zero padding d = TF 'SAME'), model.py:80-116, methods.py:58-76,116-125, gradient w.r.t. x only.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


class TorchStep:
    def __init__(self, w, cont_ids, style_ids, gatys, lambd, nb_channels=128, cnt_channels=128,
                 n_layers=None, dtype=torch.float32):
        self.cont_ids, self.style_ids = list(cont_ids), list(style_ids)
        self.gatys, self.lambd = bool(gatys), float(lambd)
        self.nb, self.cnt = nb_channels, cnt_channels
        self.NL = n_layers if n_layers is not None else min(max(self.cont_ids + self.style_ids), 29) + 1
        t = lambda a: torch.tensor(a, dtype=dtype)
        # F.conv1d weight is [Cout, Cin, K]; ours is [K, Cin, Cout]
        self.W0 = t(w.W0).permute(2, 1, 0).contiguous()
        self.b0 = t(w.b0)
        self.Wd = [t(w.Wd[l]).permute(2, 1, 0).contiguous() for l in range(self.NL)]
        self.bd = [t(w.bd[l]) for l in range(self.NL)]
        self.Wr = [t(w.Wr[l]).t().contiguous()[:, :, None] for l in range(self.NL)]
        self.br = [t(w.br[l]) for l in range(self.NL)]
        self.dtype = dtype

    def forward(self, x):
        """x [B, T] -> list of extracts (each [B, C, T]), index l = cfg.extracts[l]."""
        h = F.conv1d(x[:, None, :] / 128.0, self.W0, self.b0, padding=1)
        ex = []
        for l in range(self.NL):
            d = 1 << (l % 10)
            z = F.conv1d(torch.relu(h), self.Wd[l], self.bd[l], padding=d, dilation=d)
            h = h + F.conv1d(torch.relu(z), self.Wr[l], self.br[l])
            ex.append(h)
        return ex

    def heads(self, ex):
        g = lambda i: ex[min(i, 29)]
        Ec = torch.cat([g(i)[:, :self.cnt, :] for i in self.cont_ids], dim=1).transpose(1, 2)  # [B,T,n*cnt]
        S = torch.stack([g(i) for i in self.style_ids], dim=1)                                # [B,L,C,T]
        if not self.gatys:
            S = S.transpose(1, 2)                                                             # [B,C,L,T]
        G = S @ S.transpose(2, 3)
        ss = (G * G).sum(dim=(2, 3), keepdim=True)
        Gn = G * torch.rsqrt(torch.clamp(ss, min=1e-12))
        if self.nb < 128 and not self.gatys:
            Gn = Gn[:, :self.nb]
        return Ec, Gn

    def loss(self, x, Pc, Ps):
        Ec, Gn = self.heads(self.forward(x))
        Lc = 10.0 * ((Ec - Pc) ** 2).mean(dim=(1, 2))
        Ls = 1e3 * ((Gn - Ps) ** 2).mean(dim=(1, 2, 3))
        return Lc + self.lambd * Ls, Lc, Ls

    def step(self, x, Pc, Ps):
        """x [B,T]; Pc [B,T,n*cnt]; Ps [B,...].  Returns (losses [B,3] = total,content,style; dx [B,T])."""
        x = x.detach().clone().requires_grad_(True)
        tot, Lc, Ls = self.loss(x, Pc, Ps)
        (dx,) = torch.autograd.grad(tot.sum(), x)
        return torch.stack([tot, Lc, Ls], dim=1).detach(), dx
