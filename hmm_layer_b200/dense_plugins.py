"""Minimal dense emitter / transitioner plugins that satisfy the HmmCell plugin interface (reference
MsaHmmCell.py:43-59,68-70,80,132,145-149): used for profile-HMM shaped models (MsaHmmCell, q = 2L+3).  The
reference's own ProfileHMMEmitter / ProfileHMMTransitioner cannot be constructed as shipped; these
take the matrices directly.  `DenseEmitter.forward` is the contraction of reference Emitter.py:151-157,
`DenseTransitioner.forward` the product of Transitioner.py:320-331."""
import torch
import torch.nn as nn


class DenseTransitioner(nn.Module):
    def __init__(self, A, pi):
        """A (k, q, q) row-stochastic, pi (k, q)."""
        super().__init__()
        self.register_buffer("A_param", torch.as_tensor(A, dtype=torch.float32).clone())
        self.register_buffer("pi_param", torch.as_tensor(pi, dtype=torch.float32).clone())
        self.num_models, self.num_states = self.A_param.shape[0], self.A_param.shape[1]
        self.reverse = False
        self.A = None
        self.A_transposed = None

    def set_lengths(self, lengths):
        self.lengths = lengths

    def recurrent_init(self):
        self.A = self.A_param
        self.A_transposed = self.A_param.transpose(1, 2).contiguous()

    def make_A(self):
        return self.A_param

    def make_log_A(self):
        A = self.A_param
        return torch.where(A > 0, torch.log(torch.clamp(A, min=1e-45)), torch.full_like(A, -1000.))

    def make_initial_distribution(self):
        return self.pi_param.unsqueeze(0)            # (1, k, q)

    def forward(self, inputs):
        return torch.matmul(inputs, self.A_transposed if self.reverse else self.A)

    def reversed_view(self):
        from .gene_pred_hmm_transitioner import _ReversedTransitioner
        return _ReversedTransitioner(self)

    def get_prior_log_densities(self):
        return {"none": 0.}


class DenseEmitter(nn.Module):
    def __init__(self, B):
        """B (k, q, s): emission distribution of every state over the s input symbols."""
        super().__init__()
        self.register_buffer("B_param", torch.as_tensor(B, dtype=torch.float32).clone())
        self.num_models = self.B_param.shape[0]
        self.B = None

    def set_lengths(self, lengths):
        self.lengths = lengths

    def recurrent_init(self):
        self.B = self.B_param

    def make_B(self):
        return self.B_param

    def onehot_table(self):
        """(k, s, q): column x of B for every symbol x -- what a one-hot input row picks."""
        return self.B_param.transpose(1, 2)

    def forward(self, inputs, end_hints=None, training=False):
        """(k, b, T, s) -> (k, b, T, q): inputs . B^T (one-hot inputs pick a column of B)."""
        E = torch.einsum("kbts,kqs->kbtq", inputs, self.B_param)
        if end_hints is not None:
            E = torch.cat([E[:, :, :1] * end_hints[:, :, :1], E[:, :, 1:-1], E[:, :, -1:] * end_hints[:, :, 1:]], dim=2)
        return E

    def get_prior_log_density(self):
        return torch.zeros(self.num_models, 1, device=self.B_param.device)

    def get_aux_loss(self):
        return 0.
