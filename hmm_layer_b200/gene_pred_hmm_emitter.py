"""Gene-structure HMM emitters with the reference's class names and constructor arguments
(reference: hmm_layer/gene_pred_hmm_emitter.py).

`forward` is the stand-alone CUDA emission kernel (cell.emission_probs API parity).  Inside MsaHmmLayer the
emission tensor is never materialised: the same arithmetic runs fused in the recursion kernels.
Deviations from the as-shipped reference: inputs are read-only (D5); model i reads inputs[i] (D13)."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import kmer


def assert_codons(codons):
    assert abs(sum(p for _, p in codons) - 1) < 1e-9, "codon probabilities must sum to 1, got: " + str(codons)
    for triplet, prob in codons:
        assert len(triplet) == 3, "triplets must have length 3, got: " + str(codons)
        assert 0 <= prob <= 1, "probabilities must be in [0,1], got: " + str(codons)


def make_codon_probs(codons, pivot_left):
    assert_codons(codons)
    probs = sum(prob * kmer.encode_kmer_string(triplet, pivot_left) for triplet, prob in codons)
    return probs.reshape(64).unsqueeze(0).unsqueeze(0)


def _differentiable_emission(inputs, Bfull, codon5, num_copies, end_hints, training):
    """Emissions that stay attached to the autograd graph of B and of the class probabilities (training a model the fused
    15-state kernels do not cover).  The class contraction is the only differentiable factor; everything else (codon
    factors, end hints, the training epsilon) is a constant of the graph and comes from the CUDA emission kernel run with
    a unit class factor (class row = e_0, B = ones)."""
    from . import ops
    k, B, T, w = inputs.shape
    s = w - 5 if codon5 is not None else w
    q = Bfull.shape[1]
    with torch.no_grad():
        unit = torch.zeros_like(inputs)
        unit[..., 0] = 1.0
        if codon5 is not None:
            unit[..., s:] = inputs[..., s:]
        const = ops.emission_general(unit, torch.ones((k, q, s), dtype=torch.float32, device=inputs.device), codon5,
                                     num_copies, end_hints=end_hints, training=training)
    return torch.einsum("kbts,kqs->kbtq", inputs[..., :s], Bfull) * const


class SimpleGenePredHMMEmitter(nn.Module):
    """Class-probability emissions: emit = x_t · softmax(kernel)^T (reference :8-146)."""

    def __init__(self, num_models=1, num_copies=1, init=0., trainable_emissions=True, emit_embeddings=False,
                 embedding_dim=None, full_covariance=False, embedding_kernel_init="random_normal",
                 initial_variance=1., temperature=1., share_intron_parameters=True, **kwargs):
        super().__init__(**kwargs)
        if emit_embeddings:
            raise NotImplementedError("embedding (MVN) emissions are outside the forward/backward hot path")
        assert embedding_dim is None, "embedding_dim must not be given when emit_embeddings=False"
        self.num_models, self.num_copies = num_models, num_copies
        self.num_states = 1 + 6 * num_copies
        self.init = init
        self.trainable_emissions = trainable_emissions
        self.emit_embeddings, self.embedding_dim, self.full_covariance = False, None, full_covariance
        self.embedding_kernel_init, self.initial_variance, self.temperature = embedding_kernel_init, initial_variance, temperature
        self.share_intron_parameters = share_intron_parameters
        self.emission_kernel = None
        self.B = None
        self.built = False

    def build(self, input_shape):
        if self.built:
            return
        s = input_shape[-1]
        rows = self.num_states - 2 * self.num_copies * int(self.share_intron_parameters)
        if isinstance(self.init, torch.Tensor):
            init = self.init.detach().clone().to(torch.float32).reshape(self.num_models, rows, s)
        else:
            init = torch.full((self.num_models, rows, s), float(self.init))
        self.emission_kernel = nn.Parameter(init, requires_grad=self.trainable_emissions)
        self.built = True

    def recurrent_init(self):
        self.B = self.make_B()

    def make_B(self):
        return F.softmax(self.emission_kernel, dim=-1)

    def make_B_full(self):
        """(num_models, num_states, s): one row per state (the intron-sharing concat of reference :115-116
        applied to B)."""
        B = self.B if self.B is not None else self.make_B()
        if self.share_intron_parameters:
            nc = self.num_copies
            B = torch.cat([B[:, :1 + nc], B[:, 1:1 + nc], B[:, 1:1 + nc], B[:, 1 + nc:]], dim=1)
        return B

    def forward(self, inputs, end_hints=None, training=False):
        """(num_models, batch, T, s) -> (num_models, batch, T, num_states) on the GPU (reference :93-121): x . B^T with the
        intron rows repeated, end hints on the first / last position.  Model i reads inputs[i] (the shipped reference
        reads inputs[0] for every model, D13)."""
        from . import ops
        dev = inputs.device
        Bfull = self.make_B_full().to(dev)
        end_hints = None if end_hints is None else end_hints.to(dev)
        if torch.is_grad_enabled() and (Bfull.requires_grad or inputs.requires_grad):
            return _differentiable_emission(inputs, Bfull, None, self.num_copies, end_hints, training)
        return ops.emission_general(inputs, Bfull.detach(), None, self.num_copies, end_hints=end_hints, training=training)

    def get_prior_log_density(self):
        return torch.tensor([[0.]])

    def get_aux_loss(self):
        return 0.

    def get_config(self):
        return dict(num_models=self.num_models, num_copies=self.num_copies, init=self.init,
                    trainable_emissions=self.trainable_emissions, emit_embeddings=False, embedding_dim=None,
                    full_covariance=self.full_covariance, embedding_kernel_init=self.embedding_kernel_init,
                    initial_variance=self.initial_variance, temperature=self.temperature,
                    share_intron_parameters=self.share_intron_parameters)

    @classmethod
    def from_config(cls, config):
        return cls(**config)


class GenePredHMMEmitter(SimpleGenePredHMMEmitter):
    """Adds start/stop codon and splice-pattern constraints via soft 3-mers (reference :161-328)."""

    def __init__(self, start_codons, stop_codons, intron_begin_pattern, intron_end_pattern, l2_lambda=0.01,
                 nucleotide_kernel_init=None, trainable_nucleotides_at_exons=False, **kwargs):
        super().__init__(**kwargs)
        if trainable_nucleotides_at_exons:
            raise NotImplementedError("trainable exon nucleotide emissions are not used by any supported configuration")
        self.num_states = 1 + 14 * self.num_copies
        self.start_codons, self.stop_codons = start_codons, stop_codons
        self.intron_begin_pattern, self.intron_end_pattern = intron_begin_pattern, intron_end_pattern
        self.l2_lambda = l2_lambda
        self.nucleotide_kernel_init = nucleotide_kernel_init
        self.trainable_nucleotides_at_exons = False
        start = make_codon_probs(start_codons, True)
        stop = make_codon_probs(stop_codons, False)
        ibeg = make_codon_probs(intron_begin_pattern, True)
        iend = make_codon_probs(intron_end_pattern, False)
        anyc = make_codon_probs([("NNN", 1.)], False)
        not_stop = anyc * (stop == 0).float()
        not_stop = not_stop / not_stop.sum()
        # order: E2, START, EI0, EI1, EI2, IE0, IE1, IE2, STOP  (reference :205-217)
        left = torch.cat([anyc, start, ibeg, ibeg, ibeg, anyc, anyc, anyc, anyc], dim=1)
        right = torch.cat([not_stop, anyc, anyc, not_stop, anyc, iend, iend, iend, stop], dim=1)
        self.register_buffer("codon_probs", torch.cat([left, right], dim=0).contiguous())   # (2, 9, 64)

    def codon5(self, device):
        """(2,125,12) 5-symbol codon tables the kernels read, built once per device from the constant codon_probs."""
        from . import ops
        c = getattr(self, "_codon5", None)
        if c is None or c.device != torch.device(device):
            c = ops.prepare_codon(self.codon_probs.to(device))
            object.__setattr__(self, "_codon5", c)
        return c

    def forward(self, inputs, end_hints=None, training=False):
        """(num_models, batch, T, 15+5) -> (num_models, batch, T, num_states) emission probabilities on the GPU."""
        from . import ops
        dev = inputs.device
        Bfull = self.make_B_full().to(dev)
        end_hints = None if end_hints is None else end_hints.to(dev)
        if torch.is_grad_enabled() and (Bfull.requires_grad or inputs.requires_grad):
            return _differentiable_emission(inputs, Bfull, self.codon5(dev), self.num_copies, end_hints, training)
        Bfull = Bfull.detach()
        if self.num_copies != 1 or inputs.shape[-1] != 20:   # any class count / number of gene-model copies
            return ops.emission_general(inputs, Bfull, self.codon5(dev), self.num_copies, end_hints=end_hints, training=training)
        return ops.genepred_emission(inputs, Bfull, self.codon5(dev), end_hints=end_hints, training=training)

    def get_config(self):
        cfg = super().get_config()
        cfg.update(start_codons=self.start_codons, stop_codons=self.stop_codons,
                   intron_begin_pattern=self.intron_begin_pattern, intron_end_pattern=self.intron_end_pattern,
                   l2_lambda=self.l2_lambda, nucleotide_kernel_init=self.nucleotide_kernel_init,
                   trainable_nucleotides_at_exons=False)
        return cfg
