"""TotalProbabilityCell with the reference's interface (reference: hmm_layer/TotalProbabilityCell.py:5-63): the
log-space combine of chunk-conditional forward variables over chunk boundaries,
    new[j] = logsumexp_i(old[i] + cond[i, j]),   loglik = logsumexp_j new[j].

Inside MsaHmmLayer this role is played by the boundary scans of the fused kernels (csrc/gene15.cu, scaled
probability space with power-of-two bookkeeping); this class exists for callers that combine chunk results
themselves.  It is q*q work per chunk boundary on whatever device the tensors live on."""
import torch
import torch.nn as nn


class TotalProbabilityCell(nn.Module):
    def __init__(self, cell, reverse=False):
        super().__init__()
        self.cell = cell
        self.reverse = reverse

    @property
    def state_size(self):
        return (torch.Size([self.cell.max_num_states]), torch.Size([]))

    sate_size = state_size      # the reference spells it this way (:19)

    def make_initial_distribution(self):
        return self.cell.transitioner.make_initial_distribution()

    def forward(self, conditional_forward, states=None, training=None, init=False):
        old_forward, _ = states
        q = self.cell.max_num_states
        cond = conditional_forward.reshape(conditional_forward.shape[0], q, q)
        forward = torch.logsumexp(old_forward.unsqueeze(-1) + cond, dim=-2)
        loglik = torch.logsumexp(forward, dim=-1)
        return forward, (forward, loglik)

    def get_initial_state(self, batch_size=None, inputs=None, dtype=None):
        q = self.cell.max_num_states
        dev = self.cell.device
        dtype = dtype or torch.float32
        if self.reverse:
            return (torch.zeros(batch_size, q, dtype=dtype, device=dev), torch.zeros(batch_size, dtype=dtype, device=dev))
        init = self.make_initial_distribution().detach().to(dtype)                  # (1, k, q)
        init = init.repeat(batch_size // self.cell.num_models, 1, 1).transpose(0, 1).reshape(-1, q)
        return (torch.log(init), torch.zeros(batch_size, dtype=dtype, device=dev))
