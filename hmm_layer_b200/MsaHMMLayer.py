"""MsaHmmLayer with the reference's interface (reference: hmm_layer/MsaHMMLayer.py:8-224) on top of the B200
kernels.  The recursion drivers `_forward_recursion_impl`, `_backward_recursion_impl` and
`_state_posterior_log_probs_impl` (reference :227-521) keep their names and argument meaning; the RNN-wrapper
arguments they took are accepted and ignored because the whole scan is one fused device pass."""
import torch
import torch.nn as nn

from . import ops
from .gene_pred_hmm_emitter import GenePredHMMEmitter
from .gene_pred_hmm_transitioner import GenePredHMMTransitioner


def _is_fused_gene_cell(cell):
    """True when the cell is the 15-state Tiberius model the fused kernels are specialised for."""
    if len(cell.emitter) != 1:
        return False
    em, tr = cell.emitter[0], getattr(cell.transitioner, "base", cell.transitioner)
    return (isinstance(em, GenePredHMMEmitter) and em.num_copies == 1 and isinstance(tr, GenePredHMMTransitioner)
            and tr.k == 1 and tr.num_states == 15)


def _check_inputs(inputs, cell, parallel_factor):
    """CPU inputs are accepted for the fused gene model only (host path: outputs come back as CPU tensors, like the
    reference's); the parameters always live on the GPU."""
    if inputs.dim() != 4:
        raise ValueError(f"inputs must have shape (num_model, b, seq_len, s), got {tuple(inputs.shape)}")
    if inputs.shape[0] != cell.num_models:
        raise ValueError(f"inputs carry {inputs.shape[0]} models but the cell has {cell.num_models}")
    if parallel_factor < 1 or inputs.shape[2] % parallel_factor != 0:
        raise ValueError(f"parallel_factor={parallel_factor} must divide the sequence length {inputs.shape[2]}")
    if not inputs.is_cuda and not (_is_fused_gene_cell(cell) and cell.init_dist.is_cuda):
        raise ValueError("inputs must be CUDA tensors (CPU inputs only for the fused gene model with its cell on a GPU): "
                         "this layer has no CPU compute path")


def _gene_run(inputs, cell, end_hints, training, **want):
    em, tr = cell.emitter[0], getattr(cell.transitioner, "base", cell.transitioner)
    dev = cell.init_dist.device
    gp = getattr(cell, "_gene_params", None)   # kernel-ready parameter tensors of this parameter version (recurrent_init resets)
    if gp is None:
        gp = (tr.A.detach().contiguous(), cell.init_dist.detach().reshape(cell.num_models, -1).contiguous(),
              em.make_B_full().detach().contiguous(), em.codon5(dev))
        object.__setattr__(cell, "_gene_params", gp)
    if not inputs.is_cuda:   # host buffers: H2D / kernels / D2H pipelined inside the library, CPU tensors back
        return ops.genepred_run_host(inputs, *gp, end_hints=None if end_hints is None else end_hints.cpu(), training=training, **want)
    return ops.genepred_run(inputs, *gp, end_hints=end_hints, training=training, **want)


def _gene_loglik_diff(inputs, cell, training):
    """Differentiable log-likelihood of the fused gene model (training through the layer; the reference lets autograd
    trace its Python loop, MsaHmmCell.py:73-106).  A, pi and B stay attached to the parameters' graph; the kernels'
    closed-form gradient (ops.GenePredLoglik) supplies d loglik / d (inputs, A, pi, Bfull)."""
    em, tr = cell.emitter[0], getattr(cell.transitioner, "base", cell.transitioner)
    if _fused_params_ok(cell, em, tr):   # A, pi, B and their backward in one launch each instead of ~40 small ones
        A, pi, Bfull = ops.GeneParams.apply(tr.transition_kernel, tr.starting_distribution_kernel, em.emission_kernel)
    else:
        A, pi, Bfull = tr.A, cell.init_dist.reshape(cell.num_models, -1), em.make_B_full()
    return ops.GenePredLoglik.apply(inputs, A, pi, Bfull, em.codon5(inputs.device), None, training)


def _fused_params_ok(cell, em, tr):
    return (type(em) is GenePredHMMEmitter and em.share_intron_parameters and em.emission_kernel is not None
            and tuple(em.emission_kernel.shape) == (cell.num_models, 13, 15) and em.emission_kernel.is_cuda
            and tr.transition_kernel.numel() == 23 and tr.num_models == cell.num_models
            and tr.starting_distribution_kernel.numel() == 15)


def _priors(cell, result, return_prior):
    if not return_prior:
        return result
    extra = (cell.get_prior_log_density(), cell.get_aux_loss())
    return (*result, *extra) if isinstance(result, tuple) else (result, *extra)


def _forward_recursion_impl(inputs, cell, rnn=None, total_prob_rnn=None, end_hints=None, return_prior=False,
                            training=False, parallel_factor=1):
    """log alpha (num_model, b, seq_len, q) and log-likelihoods (num_model, b)  (reference :227-282)."""
    cell.recurrent_init()
    _check_inputs(inputs, cell, parallel_factor)
    if _is_fused_gene_cell(cell):
        r = _gene_run(inputs, cell, end_hints, training, want_post=False, want_alpha=True)
        res = (r["alpha"], r["loglik"])
    else:
        E = cell.emission_probs(inputs, end_hints=end_hints, training=training)
        r = ops.dense_run(E, cell.transitioner.A.detach(), cell.init_dist.detach().reshape(cell.num_models, -1), want_alpha=True)
        res = (r["alpha"], r["loglik"])
    return _priors(cell, res, return_prior)


def _backward_recursion_impl(inputs, cell, reverse_cell=None, rnn_backward=None, total_prob_rnn_rev=None,
                             end_hints=None, return_prior=False, training=False, parallel_factor=1):
    """log beta (num_model, b, seq_len, q)  (reference :322-381)."""
    cell.recurrent_init()
    _check_inputs(inputs, cell, parallel_factor)
    if _is_fused_gene_cell(cell):
        res = _gene_run(inputs, cell, end_hints, training, want_post=False, want_beta=True)["beta"]
    else:
        E = cell.emission_probs(inputs, end_hints=end_hints, training=training)
        res = ops.dense_run(E, cell.transitioner.A.detach(), cell.init_dist.detach().reshape(cell.num_models, -1), want_beta=True)["beta"]
    return _priors(cell, res, return_prior)


def _state_posterior_log_probs_impl(inputs, cell, reverse_cell=None, bidirectional_rnn=None, total_prob_rnn=None,
                                    total_prob_rnn_rev=None, end_hints=None, return_prior=False, training=False,
                                    no_loglik=False, parallel_factor=1):
    """log P(state q at position i | inputs), (num_model, b, seq_len, q)  (reference :422-521)."""
    cell.recurrent_init()
    _check_inputs(inputs, cell, parallel_factor)
    if _is_fused_gene_cell(cell):
        res = _gene_run(inputs, cell, end_hints, training, want_post=True, no_loglik=no_loglik)["post"]
    else:
        E = cell.emission_probs(inputs, end_hints=end_hints, training=training)
        res = ops.dense_run(E, cell.transitioner.A.detach(), cell.init_dist.detach().reshape(cell.num_models, -1),
                            want_post=True, no_loglik=no_loglik)["post"]
    return _priors(cell, res, return_prior)


class MsaHmmLayer(nn.Module):
    """Log-likelihoods and posterior state probabilities for batches of observations under a number of HMMs."""

    def __init__(self, cell, num_seqs=None, use_prior=True, sequence_weights=None, parallel_factor=1):
        super().__init__()
        self.cell = cell
        self.num_seqs = num_seqs
        self.use_prior = use_prior
        self.parallel_factor = parallel_factor
        if sequence_weights is not None:
            self.register_buffer("sequence_weights", torch.as_tensor(sequence_weights, dtype=torch.float32))
            self.register_buffer("weight_sum", torch.sum(self.sequence_weights))
        else:
            self.sequence_weights = None
            self.weight_sum = None
        self.reverse_cell = None
        self.built = False

    def build(self, input_shape=None):
        if self.built:
            return
        for em in self.cell.emitter:
            if hasattr(em, "build") and input_shape is not None and not getattr(em, "built", True):
                em.build(input_shape)
        self.reverse_cell = self.cell.make_reverse_direction_offspring()
        self.built = True

    def forward_recursion(self, inputs, end_hints=None, return_prior=False, training=False):
        self.build(inputs.shape)
        return _forward_recursion_impl(inputs, self.cell, end_hints=end_hints, return_prior=return_prior,
                                       training=training, parallel_factor=self.parallel_factor)

    def backward_recursion(self, inputs, end_hints=None, return_prior=False, training=False):
        self.build(inputs.shape)
        return _backward_recursion_impl(inputs, self.cell, self.reverse_cell, end_hints=end_hints,
                                        return_prior=return_prior, training=training, parallel_factor=self.parallel_factor)

    def state_posterior_log_probs(self, inputs, end_hints=None, return_prior=False, training=False, no_loglik=False):
        self.build(inputs.shape)
        return _state_posterior_log_probs_impl(inputs, self.cell, self.reverse_cell, end_hints=end_hints,
                                               return_prior=return_prior, training=training, no_loglik=no_loglik,
                                               parallel_factor=self.parallel_factor)

    def viterbi(self, inputs, end_hints=None, training=False, return_score=False):
        """Most probable state path (num_model, b, seq_len) int64 — the decoding the reference names (MsaHmmCell.py:13,
        `log_A_dense` :46-47) but does not ship.  Fused gene model only."""
        self.build(inputs.shape)
        cell = self.cell
        cell.recurrent_init()
        _check_inputs(inputs, cell, 1)
        if not _is_fused_gene_cell(cell) or not inputs.is_cuda:
            raise NotImplementedError("viterbi is implemented for the 15-state gene model with CUDA inputs")
        em, tr = cell.emitter[0], getattr(cell.transitioner, "base", cell.transitioner)
        path, score = ops.genepred_viterbi(inputs, tr.A.detach(), cell.init_dist.detach().reshape(cell.num_models, -1),
                                           em.make_B_full().detach(), em.codon5(inputs.device), end_hints=end_hints,
                                           training=training)
        return (path.long(), score) if return_score else path.long()

    def apply_sequence_weights(self, loglik, indices, aggregate=False):
        if self.sequence_weights is not None:
            weights = self.sequence_weights[indices]
            loglik = loglik * weights
            if aggregate:
                loglik = torch.sum(loglik, dim=1) / torch.sum(weights, dim=1)
                loglik = torch.mean(loglik)
        elif aggregate:
            loglik = torch.mean(loglik)
        return loglik

    def compute_prior(self, scaled=True):
        self.cell.recurrent_init()
        prior = self.cell.get_prior_log_density()
        return self._scale_prior(prior) if scaled else prior

    def _scale_prior(self, prior):
        if self.sequence_weights is not None:
            return prior / self.weight_sum
        if self.num_seqs is not None:
            return prior / self.num_seqs
        return prior

    def forward(self, inputs, indices=None, training=False):
        """loglik (num_model, b), aggregated loglik (), [prior (num_model), aux_loss ()]  (reference :180-208)."""
        inputs = inputs.to(torch.float32)
        self.build(inputs.shape)
        cell = self.cell
        diff = torch.is_grad_enabled() and inputs.is_cuda and (
            inputs.requires_grad or any(p.requires_grad for p in cell.parameters()))
        fused_gene = _is_fused_gene_cell(cell)
        cell.recurrent_init(fresh=diff and not (fused_gene and _fused_params_ok(
            cell, cell.emitter[0], getattr(cell.transitioner, "base", cell.transitioner))))
        _check_inputs(inputs, cell, self.parallel_factor)
        if diff and fused_gene:
            loglik = _gene_loglik_diff(inputs, cell, training)
        elif diff:   # any other cell: emissions, A and pi stay attached to the parameters' graph
            E = cell.emission_probs(inputs, training=training)
            loglik = ops.DenseLoglik.apply(E, cell.transitioner.A, cell.init_dist.reshape(cell.num_models, -1))
        elif _is_fused_gene_cell(cell):
            loglik = _gene_run(inputs, cell, None, training, want_post=False)["loglik"]
        else:
            E = cell.emission_probs(inputs, training=training)
            loglik = ops.dense_run(E, cell.transitioner.A.detach(), cell.init_dist.detach().reshape(cell.num_models, -1))["loglik"]
        if self.sequence_weights is not None:
            loglik = loglik.to(self.sequence_weights.device)
        loglik_mean = torch.squeeze(self.apply_sequence_weights(loglik, indices, aggregate=True))
        if self.use_prior:
            return loglik, loglik_mean, self._scale_prior(cell.get_prior_log_density()), cell.get_aux_loss()
        return loglik, loglik_mean

    def get_config(self):
        return dict(cell=self.cell, num_seqs=self.num_seqs, use_prior=self.use_prior,
                    sequence_weights=None if self.sequence_weights is None else self.sequence_weights.cpu().numpy(),
                    parallel_factor=self.parallel_factor)

    @classmethod
    def from_config(cls, config):
        return cls(**config)
