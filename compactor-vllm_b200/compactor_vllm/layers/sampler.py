"""Token sampler (reference: layers/sampler.py:5-27): greedy where temperature == 0, otherwise
Gumbel-max on the temperature-scaled logits.  Plain torch: a [B, vocab] elementwise pass + argmax is
library work, not part of the hot path; written sync-free so the decode step stays graph-capturable."""
import torch
from torch import nn


class Sampler(nn.Module):
    def forward(self, logits: torch.Tensor, temperatures: torch.Tensor) -> torch.Tensor:
        temps = temperatures.view(-1, 1).float()
        scaled = logits.float()
        greedy = temps == 0.0
        noise = torch.empty_like(scaled).exponential_(1).clamp_min_(1e-10).log_()
        sampled = scaled / torch.where(greedy, torch.ones_like(temps), temps) - noise
        return torch.where(greedy, scaled, sampled).argmax(dim=-1)
