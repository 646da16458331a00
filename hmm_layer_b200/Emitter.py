"""ProfileHMMEmitter with the reference's class name, constructor and methods (reference: hmm_layer/Emitter.py:7-157).

Emission matrix B (k, q, s) over the s = alphabet + 1 input symbols (last = terminal symbol), state order
LEFT_FLANK, M1..ML, I1..I(L-1), UNANNOTATED, RIGHT_FLANK, TERMINAL (reference :97-117): flank / insert / unannotated
states share the insertion distribution, match states have their own rows, the terminal state emits the terminal
symbol.  `forward` is the contraction inputs·B^T of reference :139-157 (a library GEMM; one-hot inputs pick columns).
Deviations from the as-shipped reference: B is padded to the largest state count (the shipped make_B pads
to len(lengths)+2, :124); the missing AminoAcidPrior default (:33) is replaced by "no prior" unless one is passed."""
import torch
import torch.nn as nn

from . import Initializers as initializers


class ProfileHMMEmitter(nn.Module):
    def __init__(self, emission_init=None, insertion_init=None, prior=None, frozen_insertions=True, **kwargs):
        super().__init__(**kwargs)
        emission_init = initializers.make_default_emission_init() if emission_init is None else emission_init
        insertion_init = initializers.make_default_insertion_init() if insertion_init is None else insertion_init
        self.emission_init = list(emission_init) if isinstance(emission_init, (list, tuple)) else [emission_init]
        self.insertion_init = list(insertion_init) if isinstance(insertion_init, (list, tuple)) else [insertion_init]
        self.prior = prior
        self.frozen_insertions = frozen_insertions
        self.lengths = None
        self.built = False
        self.B = self.B_transposed = None

    def set_lengths(self, lengths):
        self.lengths = list(lengths)
        self.num_models = len(self.lengths)
        assert len(self.lengths) == len(self.emission_init), \
            f"{len(self.emission_init)} emission initialisers for {len(self.lengths)} models"
        assert len(self.lengths) == len(self.insertion_init), \
            f"{len(self.insertion_init)} insertion initialisers for {len(self.lengths)} models"

    def build(self, input_shape):
        if self.built:
            return
        assert self.lengths is not None, "set_lengths must be called before build"
        s = input_shape[-1] - 1   # without the terminal symbol
        self.emission_kernel = nn.ParameterList(
            [nn.Parameter(initializers.materialize(init, (length, s))) for length, init in zip(self.lengths, self.emission_init)])
        self.insertion_kernel = nn.ParameterList(
            [nn.Parameter(initializers.materialize(init, (s,)), requires_grad=not self.frozen_insertions) for init in self.insertion_init])
        if self.prior is not None and hasattr(self.prior, "build"):
            self.prior.build()
        self.built = True

    def recurrent_init(self):
        if not self.built:
            return
        self.B = self.make_B()
        self.B_transposed = self.B.transpose(1, 2)

    def make_emission_matrix(self, i):
        return self.make_emission_matrix_from_kernels(self.emission_kernel[i], self.insertion_kernel[i], self.lengths[i])

    def make_emission_matrix_from_kernels(self, em, ins, length):
        """(2L+3, s+1): rows softmax(kernel) with a zero column for the terminal symbol; last row = that symbol."""
        s = em.shape[-1]
        rows = torch.cat([ins[None], em, ins[None].expand(length + 1, s)], dim=0)
        probs = torch.cat([torch.softmax(rows, dim=-1), torch.zeros((2 * length + 2, 1), dtype=em.dtype, device=em.device)], dim=-1)
        last = torch.zeros((1, s + 1), dtype=em.dtype, device=em.device)
        last[0, s] = 1.
        return torch.cat([probs, last], dim=0)

    def make_B(self):
        q = max(2 * l + 3 for l in self.lengths)
        mats = []
        for i in range(self.num_models):
            m = self.make_emission_matrix(i)
            mats.append(torch.nn.functional.pad(m, (0, 0, 0, q - m.shape[0])))
        return torch.stack(mats)

    def onehot_table(self):
        """(k, s + 1, q): column x of B for every symbol x -- what a one-hot input row picks (forward :139-157)."""
        B = self.B if self.B is not None else self.make_B()
        return B.transpose(1, 2)

    def make_B_amino(self):
        return self.make_B()

    def forward(self, inputs, end_hints=None, training=False):
        """(k, ..., s) -> (k, ..., q)."""
        if not self.built:
            self.build(inputs.shape)
        if self.B is None:
            self.recurrent_init()
        shape = inputs.shape
        x = inputs.reshape(shape[0], -1, shape[-1])
        emit = torch.einsum("kbs,kqs->kbq", x, self.B[..., :shape[-1]])
        return emit.reshape(shape[:-1] + (self.B.shape[1],))

    def get_aux_loss(self):
        return 0.

    def get_prior_log_density(self):
        if self.prior is None or not hasattr(self.prior, "forward"):
            return torch.zeros((self.num_models, 1), device=self.emission_kernel[0].device if self.built else None)
        return self.prior(self.make_B(), self.lengths)

    def get_config(self):
        return dict(emission_init=self.emission_init, insertion_init=self.insertion_init, prior=self.prior,
                    frozen_insertions=self.frozen_insertions)

    @classmethod
    def from_config(cls, config):
        return cls(**config)
