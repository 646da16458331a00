"""ProfileHMMTransitioner with the reference's class name, constructor and methods (reference:
hmm_layer/Transitioner.py:11-335, kernel layout :383-401, state order :429-432,470-474), rebuilt so that it can be
constructed and runs on the device.

Parameters are the reference's: one kernel per transition type of the EXPLICIT profile HMM (match / insert / delete /
begin / end / flanks, silent states included, :383-401), softmax-normalised over the out-edges of every state
(:173-195,337-380).  The kernels consume the IMPLICIT model (no silent states, q = 2L+3): delete paths are folded into
begin->match, match->match-skip and match->end probabilities (:206-258).  The construction here is vectorised torch on
whatever device the parameters live on and differentiable; no Python loop over the model length.

Deviations from the as-shipped reference: `build` accepts arrays / callables / scalars as
initialisers (the shipped one passes initialiser objects to torch.tensor, :76); out-edge probabilities are picked
with 2-D indexing (the shipped torch.gather, :186, is not a gather_nd); the match-skip block is the UPPER triangle
(target at least two match states to the right) — the shipped torch.tril mask (:221) selects the mirrored entries,
which are not probabilities; `recurrent_init` may run before `set_lengths` (MsaHmmCell.py:39 vs :179-182) and is a
no-op then.  tests/test_profile_hmm.py checks the folding against explicit silent-state elimination."""
import numpy as np
import torch
import torch.nn as nn

from . import Initializers as initializers

APPROX_LOG_ZERO = -1000.

EXPLICIT_PARTS = ["begin_to_match", "match_to_end", "match_to_match", "match_to_insert", "insert_to_match",
                  "insert_to_insert", "match_to_delete", "delete_to_match", "delete_to_delete", "left_flank_loop",
                  "left_flank_exit", "unannotated_segment_loop", "unannotated_segment_exit", "right_flank_loop",
                  "right_flank_exit", "end_to_unannotated_segment", "end_to_right_flank", "end_to_terminal"]
SHARED = {"right_flank_loop": "left_flank_loop", "right_flank_exit": "left_flank_exit"}   # reference :103-104


def explicit_part_lengths(L):
    """Kernel entries per transition type (reference :383-401)."""
    one = ["left_flank_loop", "left_flank_exit", "unannotated_segment_loop", "unannotated_segment_exit",
           "right_flank_loop", "right_flank_exit", "end_to_unannotated_segment", "end_to_right_flank", "end_to_terminal"]
    n = {"begin_to_match": L, "match_to_end": L, "match_to_delete": L, "delete_to_match": L}
    n.update({p: L - 1 for p in ("match_to_match", "match_to_insert", "insert_to_match", "insert_to_insert", "delete_to_delete")})
    n.update({p: 1 for p in one})
    return [(p, n[p]) for p in EXPLICIT_PARTS]


def explicit_source_states(L):
    """Source state of every kernel entry, part by part, in the explicit numbering LEFT_FLANK, M1..ML, I1..I(L-1),
    UNANNOTATED, RIGHT_FLANK, TERMINAL, BEGIN, END, D1..DL (reference :464-503)."""
    m = np.arange(1, L + 1)
    us, rf, begin, end, d0, i0 = 2 * L, 2 * L + 1, 2 * L + 3, 2 * L + 4, 2 * L + 5, L + 1
    src = {"begin_to_match": np.full(L, begin), "match_to_end": m, "match_to_match": m[:-1], "match_to_insert": m[:-1],
           "insert_to_match": i0 + np.arange(L - 1), "insert_to_insert": i0 + np.arange(L - 1),
           "match_to_delete": np.concatenate([[begin], m[:-1]]), "delete_to_match": d0 + np.arange(L),
           "delete_to_delete": d0 + np.arange(L - 1), "left_flank_loop": [0], "left_flank_exit": [0],
           "unannotated_segment_loop": [us], "unannotated_segment_exit": [us], "right_flank_loop": [rf],
           "right_flank_exit": [rf], "end_to_unannotated_segment": [end], "end_to_right_flank": [end], "end_to_terminal": [end]}
    return np.concatenate([np.asarray(src[p], dtype=np.int64) for p in EXPLICIT_PARTS])


def out_edge_softmax(kernel, src, num_states):
    """Probability of every kernel entry = softmax over the entries leaving the same state, with the reference's
    numerical guards (clamp at approx_log_zero + 1, +1e-16 before the final normalisation; :356,374-378)."""
    kernel = torch.clamp(kernel, min=APPROX_LOG_ZERO + 1.0)
    top = torch.full((num_states,), -float("inf"), dtype=kernel.dtype, device=kernel.device).scatter_reduce(
        0, src, kernel.detach(), reduce="amax")
    e = torch.exp(kernel - top[src])
    p = e / torch.zeros(num_states, dtype=kernel.dtype, device=kernel.device).index_add(0, src, e)[src] + 1e-16
    return p / (torch.zeros(num_states, dtype=kernel.dtype, device=kernel.device).index_add(0, src, p)[src] + 1e-16)


def fold_delete_states(lp, L):
    """Explicit log-probabilities (dict part -> tensor) -> dense (2L+3, 2L+3) transition matrix of the implicit model
    and the pieces the initial distribution needs.  Restates reference :206-258 (log space, then exp :292)."""
    t = lp["match_to_match"]
    dt, dev = t.dtype, t.device
    q, us, rf, term = 2 * L + 3, 2 * L, 2 * L + 1, 2 * L + 2
    # skip[i, j]: source i in (BEGIN, M1..M(L-1)) -> D(i+1) -> ... -> D(j+1) -> target j in (M2..ML, END), j >= i
    cum = torch.cat([torch.zeros(1, dtype=dt, device=dev), torch.cumsum(lp["delete_to_delete"], 0)])
    skip = lp["match_to_delete"][:, None] + (cum[None, :] - cum[:, None]) + lp["delete_to_match"][None, :]
    ahead = torch.ones((L, L), dtype=torch.bool, device=dev).triu()
    skip = torch.where(ahead, skip, torch.full_like(skip, APPROX_LOG_ZERO))   # j < i is no path (and would overflow in exp)
    low = torch.full((1,), APPROX_LOG_ZERO, dtype=dt, device=dev)
    entry = torch.logaddexp(lp["begin_to_match"], torch.cat([low, skip[0, :-1]]))         # BEGIN => M1..ML
    leave = torch.logaddexp(lp["match_to_end"], torch.cat([skip[1:, -1], low]))           # M1..ML => END
    through = skip[0, -1]                                                                  # BEGIN => END, all deleted
    e_us, e_rf, e_t = lp["end_to_unannotated_segment"], lp["end_to_right_flank"], lp["end_to_terminal"]
    A = torch.zeros((q, q), dtype=dt, device=dev)
    mm = torch.exp(lp["match_to_match"])
    if L > 1:
        far = torch.triu(torch.exp(skip[1:, :-1]), diagonal=1)      # M(a+1) -> M(b+2) needs b >= a+1
        A[1:L, 2:L + 1] = far + torch.diag(mm)
        ar = torch.arange(L - 1, device=dev)
        A[1 + ar, L + 1 + ar] = torch.exp(lp["match_to_insert"])
        A[L + 1 + ar, 2 + ar] = torch.exp(lp["insert_to_match"])
        A[L + 1 + ar, L + 1 + ar] = torch.exp(lp["insert_to_insert"])
    A[1:L + 1, us] = torch.exp(leave + e_us)
    A[1:L + 1, rf] = torch.exp(leave + e_rf)
    A[1:L + 1, term] = torch.exp(leave + e_t)
    for row, loop, ex in ((0, lp["left_flank_loop"], lp["left_flank_exit"]),
                          (us, lp["unannotated_segment_loop"], lp["unannotated_segment_exit"])):
        A[row, 1:L + 1] = torch.exp(ex + entry)
        A[row, us] = torch.exp(ex + through + e_us)[0]
        A[row, rf] = torch.exp(ex + through + e_rf)[0]
        A[row, term] = torch.exp(ex + through + e_t)[0]
        A[row, row] = A[row, row] + torch.exp(loop)[0]   # the unannotated loop adds to its re-entry through all deletes
    A[rf, rf] = torch.exp(lp["right_flank_loop"])[0]
    A[rf, term] = torch.exp(lp["right_flank_exit"])[0]
    A[term, term] = 1.
    return A, entry, through


class ProfileHMMTransitioner(nn.Module):
    """Default profile-HMM transition logic with the Plan7-style flank / unannotated-segment states."""

    def __init__(self, transition_init=None, flank_init=None, prior=None, frozen_kernels={}, **kwargs):
        super().__init__(**kwargs)
        transition_init = initializers.make_default_transition_init() if transition_init is None else transition_init
        flank_init = initializers.make_default_flank_init() if flank_init is None else flank_init
        self.transition_init = [transition_init] if isinstance(transition_init, dict) else list(transition_init)
        self.flank_init = list(flank_init) if isinstance(flank_init, (list, tuple)) else [flank_init]
        self.prior = prior
        self.frozen_kernels = dict(frozen_kernels)
        self.approx_log_zero = APPROX_LOG_ZERO
        self.reverse = False
        self.lengths = None
        self.built = False
        self.A = self.A_transposed = None

    def set_lengths(self, lengths):
        self.lengths = list(lengths)
        self.num_models = len(self.lengths)
        self.num_states = [2 * l + 3 for l in self.lengths]
        self.num_states_implicit = [3 * l + 5 for l in self.lengths]
        self.max_num_states = max(self.num_states)
        self.explicit_transition_kernel_parts = [explicit_part_lengths(l) for l in self.lengths]
        assert len(self.lengths) == len(self.transition_init), \
            f"{len(self.transition_init)} transition initialisers for {len(self.lengths)} models"
        assert len(self.lengths) == len(self.flank_init), f"{len(self.flank_init)} flank initialisers for {len(self.lengths)} models"
        for init, parts in zip(self.transition_init, self.explicit_transition_kernel_parts):
            names = [p for p, _ in parts]
            for p in names:
                assert p in init, "No initializer found for kernel " + p + "."
            for p in init:
                assert p in names, p + " is in the kernel init dict but there is no kernel part matching it. Wrong spelling?"
        for i, l in enumerate(self.lengths):
            self.register_buffer(f"_src{i}", torch.as_tensor(explicit_source_states(l)), persistent=False)

    def build(self, input_shape=None):
        if self.built:
            return
        assert self.lengths is not None, "set_lengths must be called before build"
        self.transition_kernel = nn.ModuleList()
        for init, parts in zip(self.transition_init, self.explicit_transition_kernel_parts):
            model = nn.ParameterDict()
            for name, n in parts:
                if name in SHARED and SHARED[name] in model:
                    continue   # right flank shares the left flank's kernels (reference :103-104)
                model[name] = nn.Parameter(initializers.materialize(init[name], (n,)),
                                           requires_grad=not self.frozen_kernels.get(name, False))
            self.transition_kernel.append(model)
        self.flank_init_kernel = nn.ParameterList([nn.Parameter(initializers.materialize(f, (1,))) for f in self.flank_init])
        if self.prior is not None and hasattr(self.prior, "build"):
            self.prior.build()
        self.built = True

    def _kernel(self, i, name):
        model = self.transition_kernel[i]
        return model[name] if name in model else model[SHARED[name]]

    def recurrent_init(self):
        if self.lengths is None:
            return
        self.build()
        self.A, self.implicit_entry, self.log_probs, self.probs = self._make_A(True)
        self.A_transposed = self.A.transpose(1, 2)
        self.A_t = self.A_transposed

    def make_transition_kernel(self):
        return [torch.cat([self._kernel(i, name) for name, _ in parts]) for i, parts in enumerate(self.explicit_transition_kernel_parts)]

    def make_probs(self):
        """Per model: dict transition type -> probabilities (reference :173-195)."""
        out = []
        for i, (parts, kernel) in enumerate(zip(self.explicit_transition_kernel_parts, self.make_transition_kernel())):
            p = out_edge_softmax(kernel, getattr(self, f"_src{i}"), self.num_states_implicit[i])
            out.append(dict(zip([n for n, _ in parts], torch.split(p, [n for _, n in parts]))))
        return out

    def make_log_probs(self):
        probs = self.make_probs()
        return [{k: torch.log(v) for k, v in m.items()} for m in probs], probs

    def _make_A(self, with_parts=False):
        self.build()
        log_probs, probs = self.make_log_probs()
        mats, entries = [], []
        for lp, L, q in zip(log_probs, self.lengths, self.num_states):
            A, entry, through = fold_delete_states(lp, L)
            pad = self.max_num_states - q
            mats.append(torch.nn.functional.pad(A, (0, pad, 0, pad)))
            entries.append((entry, through))
        A = torch.stack(mats)
        return (A, entries, log_probs, probs) if with_parts else A

    def make_A(self):
        return self._make_A()

    def make_log_A(self):
        A = self.make_A()
        return torch.where(A > 0, torch.log(torch.clamp(A, min=1e-45)), torch.full_like(A, self.approx_log_zero))

    def make_flank_init_prob(self):
        return torch.sigmoid(torch.stack([k for k in self.flank_init_kernel]))

    def make_initial_distribution(self):
        """(1, k, q): start in LEFT_FLANK with the flank probability, otherwise at BEGIN (reference :123-160)."""
        self.build()
        log_probs, _ = self.make_log_probs()
        flank = self.make_flank_init_prob()
        out = []
        for i, (lp, L, q) in enumerate(zip(log_probs, self.lengths, self.num_states)):
            _, entry, through = fold_delete_states(lp, L)
            rest = torch.log1p(-flank[i])
            d = torch.zeros(self.max_num_states, dtype=entry.dtype, device=entry.device)
            d[0] = flank[i, 0]
            d[1:L + 1] = torch.exp(rest + entry)
            d[2 * L] = torch.exp(rest + through + lp["end_to_unannotated_segment"])[0]
            d[2 * L + 1] = torch.exp(rest + through + lp["end_to_right_flank"])[0]
            d[2 * L + 2] = torch.exp(rest + through + lp["end_to_terminal"])[0]
            out.append(d)
        return torch.stack(out).unsqueeze(0)

    def forward(self, inputs):
        """(k, b, q) -> (k, b, q): x·A, or x·A^T when self.reverse (reference :320-331)."""
        return torch.matmul(inputs, self.A_transposed if self.reverse else self.A)

    def reversed_view(self):
        from .gene_pred_hmm_transitioner import _ReversedTransitioner
        return _ReversedTransitioner(self)

    def get_prior_log_densities(self):
        if self.prior is None or not hasattr(self.prior, "forward"):
            return {"none": 0.}
        return self.prior(self.make_probs(), self.make_flank_init_prob())

    def get_config(self):
        return dict(transition_init=self.transition_init, flank_init=self.flank_init, prior=self.prior,
                    frozen_kernels=self.frozen_kernels)

    @classmethod
    def from_config(cls, config):
        return cls(**config)
