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
    reference's).  The compute always runs on a GPU: a cell whose parameters live on the CPU (the reference's own script
    never moves it, tests/parallel_rnn_forward.py:42-51) has its few derived parameter tensors uploaded per call."""
    if inputs.dim() == 3 and not inputs.is_floating_point() and not _is_fused_gene_cell(cell):
        pass   # integer symbols (num_model, b, seq_len): the fused one-hot route of the dense engines
    elif inputs.dim() != 4:
        raise ValueError(f"inputs must have shape (num_model, b, seq_len, s), got {tuple(inputs.shape)}")
    if inputs.shape[0] != cell.num_models:
        raise ValueError(f"inputs carry {inputs.shape[0]} models but the cell has {cell.num_models}")
    if parallel_factor < 1 or inputs.shape[2] % parallel_factor != 0:
        raise ValueError(f"parallel_factor={parallel_factor} must divide the sequence length {inputs.shape[2]}")
    if not torch.cuda.is_available():
        raise RuntimeError("hmm_layer_b200 has no CPU compute path: a CUDA device is required")


def _compute_device(inputs, cell):
    if inputs.is_cuda:
        return inputs.device
    p = next(cell.parameters(), None)
    if p is not None and p.is_cuda:
        return p.device
    return torch.device("cuda", torch.cuda.current_device())


def _set_plain(module, name, value):
    """module.name = value for a derived quantity that is a plain attribute (not a Parameter / buffer / submodule)"""
    if name in module._parameters or name in module._buffers or name in module._modules:
        setattr(module, name, value)
    else:
        module.__dict__[name] = value


def _gene_param_tensors(cell, dev):
    """Kernel-ready (A, pi, Bfull, codon5) of the fused gene model on `dev`, rebuilt from the parameters on every call
    like the reference's recurrent_init (MsaHmmCell.py:41-50) -- one launch of the parameter kernel when the kernels
    live on the GPU -- or kept across calls after cell.freeze_derived(True)."""
    em, tr = cell.emitter[0], getattr(cell.transitioner, "base", cell.transitioner)
    if cell.__dict__.get("_frozen", False):
        cell.recurrent_init()          # resets _gene_params when a parameter changed
        gp = cell.__dict__.get("_gene_params")
        if gp is not None and gp[0].device == dev:
            return gp
    elif _fused_params_ok(cell, em, tr) and em.emission_kernel.device == dev:
        A, pi, Bfull = ops.genepred_params(tr.transition_kernel, tr.starting_distribution_kernel, em.emission_kernel)
        # what recurrent_init would have left on the plugins (values identical to make_A / make_initial_distribution);
        # plain attributes go straight into the instance dictionaries (nn.Module.__setattr__ costs 5 us apiece)
        _set_plain(tr, "A", A); _set_plain(tr, "A_transposed", A.transpose(1, 2))
        _set_plain(cell, "init_dist", pi.reshape(1, cell.num_models, -1))
        _set_plain(em, "B", None)
        return (A, pi, Bfull, em.codon5(dev))
    else:
        cell.recurrent_init()
    with torch.no_grad():
        gp = (tr.A.detach().to(dev).contiguous(), cell.init_dist.detach().reshape(cell.num_models, -1).to(dev).contiguous(),
              em.make_B_full().detach().to(dev).contiguous(), em.codon5(dev))
    object.__setattr__(cell, "_gene_params", gp)
    return gp


def _gene_run(inputs, cell, end_hints, training, **want):
    dev = _compute_device(inputs, cell)
    gp = _gene_param_tensors(cell, dev)
    if not inputs.is_cuda:   # host buffers: H2D / kernels / D2H pipelined inside the library, CPU tensors back
        with torch.cuda.device(dev):
            return ops.genepred_run_host(inputs, *gp, end_hints=None if end_hints is None else end_hints.cpu(), training=training, **want)
    if end_hints is not None:
        end_hints = end_hints.to(dev)
    return ops.genepred_run(inputs, *gp, end_hints=end_hints, training=training, **want)


def _dense_run_cell(inputs, cell, end_hints, training, **want):
    """Any other cell: emissions from the cell's emitters (CUDA kernels / cuBLAS), then the generic dense recurrence.
    Host inputs and a CPU-resident cell are accepted like in the reference: the compute runs on the GPU and the results
    come back on the inputs' device."""
    cell.recurrent_init()
    dev = _compute_device(inputs, cell)
    x = inputs.to(dev)
    table = _onehot_table(cell, x, end_hints)
    if table is not None:
        # one-hot inputs: E_t = B[:, x_t] is a column gather inside the recurrence kernels (reference Emitter.py:139-157
        # contracts inputs . B^T); neither cuBLAS nor a (k,B,T,q) emission tensor on the path
        sym = x if x.dim() == 3 else x.argmax(-1)
        r = ops.dense_run_onehot(sym.to(torch.uint8), table.to(dev), cell.transitioner.A.detach().to(dev),
                                 cell.init_dist.detach().reshape(cell.num_models, -1).to(dev), **want)
        return r if inputs.is_cuda else {k: v.to(inputs.device) for k, v in r.items()}
    E = cell.emission_probs(x, end_hints=None if end_hints is None else end_hints.to(dev), training=training)
    r = ops.dense_run(E.detach(), cell.transitioner.A.detach().to(dev), cell.init_dist.detach().reshape(cell.num_models, -1).to(dev), **want)
    return r if inputs.is_cuda else {k: v.to(inputs.device) for k, v in r.items()}


def _onehot_table(cell, x, end_hints):
    """(k, s, q) emission table for the fused one-hot route, or None: taken when the caller declared the inputs one-hot
    (`cell.onehot_inputs = True`, or passes integer symbols (k, B, T) directly) and the cell's single emitter exposes
    `onehot_table()` (ProfileHMMEmitter, DenseEmitter).  Soft inputs keep the GEMM form."""
    symbols = x.dim() == 3 and not x.is_floating_point()
    if not (symbols or getattr(cell, "onehot_inputs", False)) or end_hints is not None or len(cell.emitter) != 1:
        return None
    em = cell.emitter[0]
    if not hasattr(em, "onehot_table"):
        return None
    t = em.onehot_table().detach()
    if not symbols and t.shape[1] != x.shape[-1]:
        t = t[:, :x.shape[-1]]
    return t.contiguous()


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
    _check_inputs(inputs, cell, parallel_factor)
    if _is_fused_gene_cell(cell):
        r = _gene_run(inputs, cell, end_hints, training, want_post=False, want_alpha=True)
        res = (r["alpha"], r["loglik"])
    else:
        r = _dense_run_cell(inputs, cell, end_hints, training, want_alpha=True)
        res = (r["alpha"], r["loglik"])
    return _priors(cell, res, return_prior)


def _backward_recursion_impl(inputs, cell, reverse_cell=None, rnn_backward=None, total_prob_rnn_rev=None,
                             end_hints=None, return_prior=False, training=False, parallel_factor=1):
    """log beta (num_model, b, seq_len, q)  (reference :322-381)."""
    _check_inputs(inputs, cell, parallel_factor)
    if _is_fused_gene_cell(cell):
        res = _gene_run(inputs, cell, end_hints, training, want_post=False, want_beta=True)["beta"]
    else:
        res = _dense_run_cell(inputs, cell, end_hints, training, want_beta=True)["beta"]
    return _priors(cell, res, return_prior)


def _state_posterior_log_probs_impl(inputs, cell, reverse_cell=None, bidirectional_rnn=None, total_prob_rnn=None,
                                    total_prob_rnn_rev=None, end_hints=None, return_prior=False, training=False,
                                    no_loglik=False, parallel_factor=1):
    """log P(state q at position i | inputs), (num_model, b, seq_len, q)  (reference :422-521)."""
    _check_inputs(inputs, cell, parallel_factor)
    if _is_fused_gene_cell(cell):
        res = _gene_run(inputs, cell, end_hints, training, want_post=True, no_loglik=no_loglik)["post"]
    else:
        res = _dense_run_cell(inputs, cell, end_hints, training, want_post=True, no_loglik=no_loglik)["post"]
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
        _check_inputs(inputs, cell, 1)
        if not _is_fused_gene_cell(cell) or not inputs.is_cuda:
            raise NotImplementedError("viterbi is implemented for the 15-state gene model with CUDA inputs")
        A, pi, Bfull, codon5 = _gene_param_tensors(cell, inputs.device)
        path, score = ops.genepred_viterbi(inputs, A, pi, Bfull, codon5, end_hints=end_hints, training=training)
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
        symbols = inputs.dim() == 3 and not inputs.is_floating_point()    # integer symbols: the fused one-hot route (inference)
        if not symbols:
            inputs = inputs.to(torch.float32)
        self.build(None if symbols else inputs.shape)
        cell = self.cell
        diff = not symbols and torch.is_grad_enabled() and (inputs.requires_grad or any(p.requires_grad for p in cell.parameters()))
        fused_gene = _is_fused_gene_cell(cell)
        _check_inputs(inputs, cell, self.parallel_factor)
        home = inputs.device
        if diff and not inputs.is_cuda:
            # the reference traces its loop wherever the tensors live; here the differentiable route is the CUDA one, so
            # host inputs follow the parameters to their device (a differentiable copy) instead of silently losing grad_fn
            p0 = next(cell.parameters(), None)
            if p0 is None or not p0.is_cuda:
                raise ValueError("a differentiable call needs the cell's parameters on a CUDA device (cell.to('cuda')): "
                                 "this layer has no CPU compute path")
            inputs = inputs.to(p0.device)
        if diff and fused_gene:
            if not _fused_params_ok(cell, cell.emitter[0], getattr(cell.transitioner, "base", cell.transitioner)):
                cell.recurrent_init(fresh=True)     # a fresh autograd graph from the parameters, like the reference
            loglik = _gene_loglik_diff(inputs, cell, training)
        elif diff:   # any other cell: emissions, A and pi stay attached to the parameters' graph
            cell.recurrent_init(fresh=True)
            E = cell.emission_probs(inputs, training=training)
            loglik = ops.DenseLoglik.apply(E, cell.transitioner.A, cell.init_dist.reshape(cell.num_models, -1))
        elif fused_gene:
            loglik = _gene_run(inputs, cell, None, training, want_post=False)["loglik"]
        else:
            loglik = _dense_run_cell(inputs, cell, None, training)["loglik"]
        loglik = loglik.to(home)      # results live where the inputs live, like the reference's (a differentiable copy)
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
