"""Gene-structure HMM transitioners with the reference's class names and constructor arguments
(reference: hmm_layer/gene_pred_hmm_transitioner.py).  Parameter -> A / pi construction is small host-side
PyTorch that runs on whatever device the parameters live on; the recursion kernels consume A and pi.

Deliberate deviations from the as-shipped reference:
  D6  edges whose kernel value is exactly 0.0 stay edges (the reference's dense==0 test drops them);
  D1  `reverse` is not a flag shared between the forward and the reverse cell: a reverse cell holds a
      `reversed_view()` of the transitioner that shares the parameters;
  D10 make_log_A returns log(A) (the reference applies log_softmax to probabilities);
  D11 for k>1 the kernel is initialised with the k-copy edge classification."""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


def make_transition_matrix_from_indices(indices, kernel, num_states, approx_log_zero=-1000.):
    """Dense (q,q) row-stochastic matrix: softmax of the kernel over each state's out-edges
    (reference Transitioner.py:337-380, mask taken from the edge list)."""
    idx = torch.as_tensor(np.asarray(indices), device=kernel.device)
    kernel = torch.clamp(kernel, min=approx_log_zero + 1.0)
    dense = torch.full((num_states, num_states), approx_log_zero, dtype=kernel.dtype, device=kernel.device)
    dense = dense.index_put((idx[:, 0], idx[:, 1]), kernel)
    mask = torch.zeros_like(dense).index_put((idx[:, 0], idx[:, 1]), torch.ones_like(kernel))
    probs = (torch.softmax(dense, -1) + 1e-16) * mask
    return probs / (probs.sum(-1, keepdim=True) + 1e-16)


class SimpleGenePredHMMTransitioner(nn.Module):
    """7 states Ir, I0-2, E0-2 / 15 edges (reference :8-182)."""

    def __init__(self, num_models=1, initial_exon_len=100, initial_intron_len=10000, initial_ir_len=10000, init=None,
                 starting_distribution_init="zeros", starting_distribution_trainable=True,
                 transitions_trainable=True, init_component_sd=0, **kwargs):
        super().__init__(**kwargs)
        self.num_models = num_models
        self.initial_exon_len, self.initial_intron_len, self.initial_ir_len = initial_exon_len, initial_intron_len, initial_ir_len
        if not hasattr(self, "num_states"):
            self.num_states = 7
        if not hasattr(self, "k"):
            self.k = 1
        self.indices = self.make_transition_indices()
        self.starting_distribution_init = starting_distribution_init
        self.starting_distribution_trainable = starting_distribution_trainable
        self.transitions_trainable = transitions_trainable
        self.num_transitions = len(self.indices)
        self.reverse = False
        self.init = self.make_transition_init(self.k, init_component_sd) if init is None else np.asarray(init, dtype=np.float64)
        self.transition_kernel = nn.Parameter(torch.tensor(self.init, dtype=torch.float32).reshape(1, -1),
                                              requires_grad=transitions_trainable)
        start = torch.zeros(1, 1, self.num_states) if starting_distribution_init == "zeros" else torch.ones(1, 1, self.num_states)
        self.starting_distribution_kernel = nn.Parameter(start, requires_grad=starting_distribution_trainable)
        self.A = None
        self.A_transposed = None

    # ---- edge classification used by the initialiser (reference :46-64) ----
    def is_intergenic_loop(self, e): return e[1] == e[2] == 0
    def is_intron_loop(self, e, k=1): return e[1] == e[2] and 0 < e[1] < 1 + 3 * k
    def is_exon_transition(self, e, k=1):
        off = 1 + 3 * k
        return off <= e[1] < off + 3 * k and e[2] - off == (e[1] - off + k) % (3 * k)
    def is_exon_1_out_transition(self, e, k=1): return 1 + 4 * k <= e[1] < 1 + 5 * k and e[1] != e[2]
    def is_intergenic_out_transition(self, e, k=1): return e[1] == 0 and e[2] != 0

    def make_transition_indices(self, model_index=0):
        Ir, I, E = 0, [1, 2, 3], [4, 5, 6]
        idx = [(Ir, Ir), (Ir, E[0]), (E[2], Ir)]
        for c in range(3):
            idx += [(E[c], E[(c + 1) % 3]), (E[c], I[c]), (I[c], I[c]), (I[c], E[(c + 1) % 3])]
        return np.concatenate([np.full((len(idx), 1), model_index, dtype=np.int64), np.array(idx, dtype=np.int64)], 1)

    def make_transition_init(self, k=1, sd=0.05):
        def logit(p): return -np.log(1 / p - 1)
        out = []
        for e in self.indices:
            if self.is_intergenic_loop(e): out.append(logit(1 - 1. / self.initial_ir_len))
            elif self.is_intron_loop(e, k): out.append(logit(1 - 1. / self.initial_intron_len))
            elif self.is_exon_transition(e, k): out.append(logit(1 - 1. / self.initial_exon_len))
            elif self.is_exon_1_out_transition(e, k): out.append(np.log(1. / 2))
            elif self.is_intergenic_out_transition(e, k): out.append(np.log(1. / k) + np.random.normal(0., sd))
            else: out.append(0.)
        return np.array(out)

    def recurrent_init(self):
        self.A = self.make_A()
        self.A_transposed = self.A.transpose(1, 2)

    def make_A(self):
        A = make_transition_matrix_from_indices(self.indices[:, 1:], self.transition_kernel.reshape(-1), self.num_states)
        return A.unsqueeze(0).repeat(self.num_models, 1, 1)

    def make_A_sparse(self, values=None):
        v = self.transition_kernel.reshape(-1) if values is None else values
        A = make_transition_matrix_from_indices(self.indices[:, 1:], v, self.num_states)
        return A.unsqueeze(0).to_sparse()

    def make_log_A(self):
        A = self.make_A()
        return torch.where(A > 0, torch.log(torch.clamp(A, min=1e-45)), torch.full_like(A, -1000.))

    def make_initial_distribution(self):
        return F.softmax(self.starting_distribution_kernel, -1).repeat(1, self.num_models, 1)

    def forward(self, inputs):
        """(k, b, q) -> (k, b, q): x·A, or x·A^T when self.reverse (reference :114-125)."""
        return torch.matmul(inputs, self.A_transposed if self.reverse else self.A)

    def reversed_view(self):
        return _ReversedTransitioner(self)

    def get_prior_log_densities(self):
        return {"none": 0.}

    def get_config(self):
        return dict(initial_exon_len=self.initial_exon_len, initial_intron_len=self.initial_intron_len,
                    initial_ir_len=self.initial_ir_len, starting_distribution_init=self.starting_distribution_init,
                    starting_distribution_trainable=self.starting_distribution_trainable,
                    transitions_trainable=self.transitions_trainable)

    @classmethod
    def from_config(cls, config):
        return cls(**config)


class _ReversedTransitioner(nn.Module):
    """Shares every parameter/cache with `base` but multiplies by A^T."""
    reverse = True

    def __init__(self, base):
        super().__init__()
        object.__setattr__(self, "base", base)      # not registered as a sub-module: parameters stay owned by base

    def __getattr__(self, name):
        if name == "base":
            return object.__getattribute__(self, "base")
        return getattr(object.__getattribute__(self, "base"), name)

    def recurrent_init(self): self.base.recurrent_init()
    def forward(self, inputs): return torch.matmul(inputs, self.base.A_transposed)
    def reversed_view(self): return self


class GenePredHMMTransitioner(SimpleGenePredHMMTransitioner):
    """15 states / 23 edges: Ir, I0-2, E0-2, START, EI0-2, IE0-2, STOP (reference :184-261)."""

    def __init__(self, use_experimental_prior=False, **kwargs):
        if not hasattr(self, "num_states"):
            self.num_states = 15
        if not hasattr(self, "k"):
            self.k = 1
        super().__init__(**kwargs)
        if use_experimental_prior:
            raise NotImplementedError("experimental Dirichlet prior is a training regulariser outside the forward/backward path")
        self.use_experimental_prior = False

    def make_transition_indices(self, model_index=0):
        k = self.k
        Ir = 0
        I = list(range(1, 1 + 3 * k)); E = list(range(1 + 3 * k, 1 + 6 * k))
        START = list(range(1 + 6 * k, 1 + 7 * k)); EI = list(range(1 + 7 * k, 1 + 10 * k))
        IE = list(range(1 + 10 * k, 1 + 13 * k)); STOP = list(range(1 + 13 * k, 1 + 14 * k))
        idx = [(Ir, Ir)]
        for h in range(k):
            idx += [(Ir, START[h]), (STOP[h], Ir), (START[h], E[k + h]), (E[k + h], STOP[h])]
            for c in range(3):
                idx += [(E[k * c + h], E[k * ((c + 1) % 3) + h]), (E[k * c + h], EI[k * c + h]),
                        (EI[k * c + h], I[k * c + h]), (I[k * c + h], I[k * c + h]),
                        (I[k * c + h], IE[k * c + h]), (IE[k * c + h], E[k * c + h])]
        assert len(idx) == 1 + 22 * k
        return np.concatenate([np.full((len(idx), 1), model_index, dtype=np.int64), np.array(idx, dtype=np.int64)], 1)


class GenePredMultiHMMTransitioner(GenePredHMMTransitioner):
    """k gene-model copies sharing the Ir state: 1+14k states, 1+22k edges (reference :263-308)."""

    def __init__(self, k=1, init_component_sd=0.2, **kwargs):
        self.k = k
        self.num_states = 1 + 14 * k
        self.init_component_sd = init_component_sd
        if k > 1:
            kwargs.setdefault("init_component_sd", init_component_sd)
        super().__init__(**kwargs)

    def get_config(self):
        cfg = super().get_config()
        cfg.update(k=self.k)
        return cfg
