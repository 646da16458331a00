"""Tensor-level entry points over the C ABI: validate, allocate outputs/workspace as torch CUDA tensors, pass raw
device pointers and the current CUDA stream.  PyTorch is plumbing here (memory + streams), not compute."""
import torch

from . import _lib

TRAINING = 1
NO_LOGLIK = 2
ALPHA_SCRATCH = 4
BETA_SCRATCH = 8
ALPHA_NORMALIZED = 16


def _ptr(t):
    return None if t is None else t.data_ptr()


def _req(t, name, shape=None):
    if not isinstance(t, torch.Tensor):
        raise ValueError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise ValueError(f"{name} must live on a CUDA device (no CPU fallback)")
    if t.dtype != torch.float32:
        raise ValueError(f"{name} must be float32, got {t.dtype}")
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError(f"{name} must have shape {tuple(shape)}, got {tuple(t.shape)}")
    return t.contiguous()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream(device=None):
    """cudaStream_t of torch's current stream (the raw getter skips the Stream object: 2 us instead of 14 per call)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device() if device is None or device.index is None else device.index)
    return torch.cuda.current_stream(device).cuda_stream


_ws_cache = {}


def _workspace(nbytes, device):
    key = (device.index, _stream(device))
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def set_tuning(chunk_len=0, finalize_threads=0, dense_path=0):
    """chunk_len: positions per chunk (0 = chosen per call from the problem size and the SM count);
    finalize_threads: threads per CTA of the finalize kernel (0 = derived from chunk_len);
    dense_path: step kernel of the generic dense path, 0 = by size, 1 = CUDA cores, 2 = tcgen05 tensor cores."""
    _lib.check(_lib.load().hmm_b200_set_tuning(chunk_len, finalize_threads, dense_path))


def set_fused(on=True):
    """Single-launch forward+posterior kernel for posterior-only gene-HMM calls (default on); off = two-kernel path."""
    _lib.check(_lib.load().hmm_b200_set_fused_mode(0 if on else 1))


def profile_enable(on=True):
    _lib.check(_lib.load().hmm_b200_profile_enable(1 if on else 0))


def profile_read():
    """Durations (ms) of the kernels of the last profiled call, in launch order."""
    import ctypes
    buf = (ctypes.c_float * 8)()
    n = _lib.load().hmm_b200_profile_read(buf, 8)
    if n < 0:
        _lib.check(n)
    return [buf[i] for i in range(n)]


def prepare_codon(codon):
    """(2,9,64) codon_probs -> (2,125,12) 5-symbol tables the kernels read (once per parameter set)."""
    codon = _req(codon, "codon", (2, 9, 64))
    with torch.cuda.device(codon.device):
        out = torch.empty((2, 125, 12), dtype=torch.float32, device=codon.device)
        _lib.check(_lib.load().hmm_b200_genepred_prepare_codon(_ptr(codon), _ptr(out), _stream()))
    return out


def _codon5(codon):
    if tuple(codon.shape) == (2, 125, 12):
        return _req(codon, "codon5", (2, 125, 12))
    return prepare_codon(codon)


def genepred_run(inputs, A, pi, Bfull, codon, end_hints=None, training=False, no_loglik=False,
                 want_post=True, want_alpha=False, want_beta=False, out=None, alpha_normalized=False):
    """Fused 15-state gene-HMM evaluation.  inputs (k,B,T,20); codon = the emitter's (2,9,64) codon_probs or the
    prepared (2,125,12) tables (prepare_codon).  Returns dict with the requested tensors
    ('post','alpha','beta' each (k,B,T,15)) and 'loglik' (k,B)."""
    lib = _lib.load()
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T, 20), got {tuple(inputs.shape)}")
    k, B, T, _ = inputs.shape
    inputs = _req(inputs, "inputs")
    dev = inputs.device
    A = _req(A, "A", (k, 15, 15)); pi = _req(pi.reshape(k, 15), "pi", (k, 15))
    Bfull = _req(Bfull, "Bfull", (k, 15, 15)); codon = _codon5(codon)
    if end_hints is not None:
        end_hints = _req(end_hints, "end_hints", (k, B, 2, 15))
    with torch.cuda.device(dev):
        res = {}
        out = out or {}
        for name, want in (("post", want_post), ("alpha", want_alpha), ("beta", want_beta)):
            if want:
                res[name] = out.get(name) if out.get(name) is not None else torch.empty((k, B, T, 15), dtype=torch.float32, device=dev)
        res["loglik"] = torch.empty((k, B), dtype=torch.float32, device=dev)
        nbytes = lib.hmm_b200_genepred_workspace_size(k, B, T)
        ws = _workspace(nbytes, dev)
        flags = (TRAINING if training else 0) | (NO_LOGLIK if no_loglik else 0) | (ALPHA_NORMALIZED if alpha_normalized else 0)
        _lib.check(lib.hmm_b200_genepred_run(
            _ptr(inputs), k, B, T, _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints), flags,
            _ptr(res.get("post")), _ptr(res.get("alpha")), _ptr(res.get("beta")), _ptr(res["loglik"]),
            _ptr(ws), ws.numel(), _stream()))
    return res


def pack_compact(inputs):
    """(k,B,T,20) fp32 reference layout -> (class probabilities (k,B,T,15) bf16, nucleotide codes (k,B,T) uint8:
    0..3 = A,C,G,T, 4 = N) — the compact wire format of hmm_b200_genepred_run_compact.  Works on CPU or CUDA tensors;
    rows that are neither one-hot nor N cannot be expressed and raise."""
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T, 20), got {tuple(inputs.shape)}")
    nuc = inputs[..., 15:]
    hard = ((nuc == 0) | (nuc == 1)).all(-1) & (nuc.sum(-1) == 1)
    if not bool(hard.all()):
        raise ValueError("the compact format needs one-hot (or N) nucleotide rows")
    return inputs[..., :15].to(torch.bfloat16).contiguous(), nuc.argmax(-1).to(torch.uint8).contiguous()


def genepred_run_compact(class_probs, nuc_codes, A, pi, Bfull, codon, end_hints=None, training=False, no_loglik=False,
                         want_post=True, want_alpha=False, want_beta=False):
    """The fused evaluation from the compact wire format (bf16 class probabilities (k,B,T,15) + uint8 nucleotide codes
    (k,B,T)).  CUDA tensors: device path; CPU tensors (ideally pinned): pipelined host path, pinned CPU results."""
    lib = _lib.load()
    if class_probs.dim() != 4 or class_probs.shape[-1] != 15 or class_probs.dtype != torch.bfloat16:
        raise ValueError("class_probs must be a bfloat16 tensor of shape (num_models, batch, T, 15)")
    k, B, T, _ = class_probs.shape
    if tuple(nuc_codes.shape) != (k, B, T) or nuc_codes.dtype != torch.uint8 or nuc_codes.device != class_probs.device:
        raise ValueError("nuc_codes must be a uint8 tensor of shape (num_models, batch, T) on the same device")
    class_probs, nuc_codes = class_probs.contiguous(), nuc_codes.contiguous()
    A = _req(A, "A", (k, 15, 15)); pi = _req(pi.reshape(k, 15), "pi", (k, 15))
    Bfull = _req(Bfull, "Bfull", (k, 15, 15)); codon = _codon5(codon)
    dev, host = A.device, not class_probs.is_cuda
    if end_hints is not None:
        end_hints = end_hints.to(torch.float32).contiguous() if host else _req(end_hints, "end_hints", (k, B, 2, 15))
    flags = (TRAINING if training else 0) | (NO_LOGLIK if no_loglik else 0)
    with torch.cuda.device(dev):
        kw = dict(dtype=torch.float32, pin_memory=True) if host else dict(dtype=torch.float32, device=dev)
        res = {}
        for name, want in (("post", want_post), ("alpha", want_alpha), ("beta", want_beta)):
            if want:
                res[name] = torch.empty((k, B, T, 15), **kw)
        res["loglik"] = torch.empty((k, B), **kw)
        if host:
            ws = _workspace(lib.hmm_b200_genepred_host_scratch_size(k, B, T), dev)
            fn = lib.hmm_b200_genepred_run_host_compact
        else:
            ws = _workspace(lib.hmm_b200_genepred_workspace_size(k, B, T), dev)
            fn = lib.hmm_b200_genepred_run_compact
        _lib.check(fn(_ptr(class_probs), _ptr(nuc_codes), k, B, T, _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints),
                      flags, _ptr(res.get("post")), _ptr(res.get("alpha")), _ptr(res.get("beta")), _ptr(res["loglik"]),
                      _ptr(ws), ws.numel(), _stream()))
        if host:
            torch.cuda.current_stream(dev).synchronize()
    return res


def genepred_viterbi(inputs, A, pi, Bfull, codon, end_hints=None, training=False):
    """Most probable state path of the 15-state gene HMM (hmm_b200_genepred_viterbi).  inputs (k,B,T,20) CUDA.
    Returns (path (k,B,T) uint8, score (k,B) float32 = log joint probability of the path)."""
    lib = _lib.load()
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T, 20), got {tuple(inputs.shape)}")
    k, B, T, _ = inputs.shape
    inputs = _req(inputs, "inputs")
    dev = inputs.device
    A = _req(A, "A", (k, 15, 15)); pi = _req(pi.reshape(k, 15), "pi", (k, 15))
    Bfull = _req(Bfull, "Bfull", (k, 15, 15)); codon = _codon5(codon)
    if end_hints is not None:
        end_hints = _req(end_hints, "end_hints", (k, B, 2, 15))
    with torch.cuda.device(dev):
        path = torch.empty((k, B, T), dtype=torch.uint8, device=dev)
        score = torch.empty((k, B), dtype=torch.float32, device=dev)
        ws = _workspace(lib.hmm_b200_genepred_viterbi_workspace_size(k, B, T), dev)
        _lib.check(lib.hmm_b200_genepred_viterbi(
            _ptr(inputs), k, B, T, _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints), TRAINING if training else 0,
            _ptr(path), _ptr(score), _ptr(ws), ws.numel(), _stream()))
    return path, score


def genepred_grad(inputs, A, pi, Bfull, codon, log_post, log_alpha_hat, grad_loglik, end_hints=None, training=False,
                  want_inputs=True, want_A=True, want_pi=True, want_B=True):
    """Gradient of sum(grad_loglik * loglik) of the fused gene-HMM evaluation (hmm_b200_genepred_grad).
    log_post / log_alpha_hat: 'post' and 'alpha' of genepred_run(..., want_alpha=True, alpha_normalized=True).
    Returns dict with 'inputs' (k,B,T,20), 'A' (k,15,15), 'pi' (k,15), 'Bfull' (k,15,15) as requested."""
    lib = _lib.load()
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T, 20), got {tuple(inputs.shape)}")
    k, B, T, _ = inputs.shape
    inputs = _req(inputs, "inputs")
    dev = inputs.device
    A = _req(A, "A", (k, 15, 15)); pi = _req(pi.reshape(k, 15), "pi", (k, 15))
    Bfull = _req(Bfull, "Bfull", (k, 15, 15)); codon = _codon5(codon)
    log_post = _req(log_post, "log_post", (k, B, T, 15)); log_alpha_hat = _req(log_alpha_hat, "log_alpha_hat", (k, B, T, 15))
    grad_loglik = _req(grad_loglik, "grad_loglik", (k, B))
    if end_hints is not None:
        end_hints = _req(end_hints, "end_hints", (k, B, 2, 15))
    if not (want_inputs or want_A or want_pi or want_B):
        return {}
    with torch.cuda.device(dev):
        res = {}
        if want_inputs:
            res["inputs"] = torch.empty((k, B, T, 20), dtype=torch.float32, device=dev)
        if want_A:
            res["A"] = torch.empty((k, 15, 15), dtype=torch.float32, device=dev)
        if want_pi:
            res["pi"] = torch.empty((k, 15), dtype=torch.float32, device=dev)
        if want_B:
            res["Bfull"] = torch.empty((k, 15, 15), dtype=torch.float32, device=dev)
        ws = _workspace(lib.hmm_b200_genepred_grad_workspace_size(k, B, T), dev)
        _lib.check(lib.hmm_b200_genepred_grad(
            _ptr(inputs), k, B, T, _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints), TRAINING if training else 0,
            _ptr(log_post), _ptr(log_alpha_hat), _ptr(grad_loglik), _ptr(res.get("inputs")), _ptr(res.get("A")),
            _ptr(res.get("pi")), _ptr(res.get("Bfull")), _ptr(ws), ws.numel(), _stream()))
    return res


def genepred_params(tk, sk, ek):
    """A (k,15,15), pi (k,15), Bfull (k,15,15) from the kernels in one launch, no autograd bookkeeping (the inference path
    of the layer: the parameters are rebuilt on every call like the reference's recurrent_init, so the Python in front of
    the first upload counts).  GeneParams below is the differentiable form."""
    lib = _lib.load()
    k, dev = ek.shape[0], ek.device
    A = torch.empty((k, 15, 15), dtype=torch.float32, device=dev)
    pi = torch.empty((k, 15), dtype=torch.float32, device=dev)
    Bfull = torch.empty((k, 15, 15), dtype=torch.float32, device=dev)
    tk, sk, ek = tk.detach(), sk.detach(), ek.detach()
    if not (tk.is_contiguous() and sk.is_contiguous() and ek.is_contiguous() and tk.dtype == sk.dtype == ek.dtype == torch.float32
            and tk.device == sk.device == dev):
        tk, sk, ek = _req(tk.to(dev), "transition_kernel"), _req(sk.to(dev), "start_kernel"), _req(ek, "emission_kernel")
    with torch.cuda.device(dev):
        _lib.check(lib.hmm_b200_genepred_params(tk.data_ptr(), sk.data_ptr(), ek.data_ptr(), k, A.data_ptr(), pi.data_ptr(),
                                                Bfull.data_ptr(), _stream(dev)))
    return A, pi, Bfull


class GeneParams(torch.autograd.Function):
    """(transition kernel (1,23), starting distribution kernel (1,1,15), emission kernel (k,13,15)) -> A (k,15,15),
    pi (k,15), Bfull (k,15,15) of the 15-state gene model in one launch, with its backward in one launch
    (hmm_b200_genepred_params / _params_grad).  Same arithmetic as GenePredHMMTransitioner.make_A /
    make_initial_distribution and SimpleGenePredHMMEmitter.make_B_full, which stay the reference-shaped API."""

    @staticmethod
    def forward(ctx, tk, sk, ek):
        lib = _lib.load()
        tk, sk, ek = _req(tk, "transition_kernel"), _req(sk, "start_kernel"), _req(ek, "emission_kernel")
        k = ek.shape[0]
        if tk.numel() != 23 or sk.numel() != 15 or tuple(ek.shape) != (k, 13, 15):
            raise ValueError("GeneParams: kernels of the single-copy 15-state gene model expected (23 / 15 / (k,13,15))")
        dev = ek.device
        with torch.cuda.device(dev):
            A = torch.empty((k, 15, 15), dtype=torch.float32, device=dev)
            pi = torch.empty((k, 15), dtype=torch.float32, device=dev)
            Bfull = torch.empty((k, 15, 15), dtype=torch.float32, device=dev)
            _lib.check(lib.hmm_b200_genepred_params(_ptr(tk), _ptr(sk), _ptr(ek), k, _ptr(A), _ptr(pi), _ptr(Bfull), _stream()))
        ctx.save_for_backward(tk, sk, ek)
        ctx.shapes = (tk.shape, sk.shape)
        return A, pi, Bfull

    @staticmethod
    def backward(ctx, gA, gpi, gB):
        lib = _lib.load()
        tk, sk, ek = ctx.saved_tensors
        k, dev = ek.shape[0], ek.device
        need = ctx.needs_input_grad
        c = lambda g: None if g is None else g.contiguous().to(torch.float32)
        gA, gpi, gB = c(gA), c(gpi), c(gB)
        with torch.cuda.device(dev):
            gtk = torch.empty(23, dtype=torch.float32, device=dev) if need[0] else None
            gsk = torch.empty(15, dtype=torch.float32, device=dev) if need[1] else None
            gek = torch.empty((k, 13, 15), dtype=torch.float32, device=dev) if need[2] else None
            if need[0] or need[1] or need[2]:
                _lib.check(lib.hmm_b200_genepred_params_grad(_ptr(tk), _ptr(sk), _ptr(ek), k, _ptr(gA), _ptr(gpi), _ptr(gB),
                                                             _ptr(gtk), _ptr(gsk), _ptr(gek), _stream()))
        return (None if gtk is None else gtk.reshape(ctx.shapes[0]), None if gsk is None else gsk.reshape(ctx.shapes[1]), gek)


class GenePredLoglik(torch.autograd.Function):
    """loglik (k,B) of the fused gene HMM, differentiable w.r.t. inputs (class probabilities), A, pi and Bfull.
    Forward = one fused evaluation that also keeps log posterior and normalised forward vectors; backward = the
    closed-form expected-count kernel.  codon / end_hints / training are constants of the graph."""

    @staticmethod
    def forward(ctx, inputs, A, pi, Bfull, codon, end_hints, training):
        r = genepred_run(inputs, A, pi, Bfull, codon, end_hints=end_hints, training=training, want_post=True,
                         want_alpha=True, alpha_normalized=True)
        ctx.save_for_backward(inputs, A, pi, Bfull, codon, r["post"], r["alpha"])
        ctx.end_hints, ctx.training = end_hints, training
        return r["loglik"]

    @staticmethod
    def backward(ctx, grad_loglik):
        inputs, A, pi, Bfull, codon, post, ahat = ctx.saved_tensors
        need = ctx.needs_input_grad
        g = genepred_grad(inputs, A, pi, Bfull, codon, post, ahat, grad_loglik.contiguous().to(torch.float32),
                          end_hints=ctx.end_hints, training=ctx.training,
                          want_inputs=need[0], want_A=need[1], want_pi=need[2], want_B=need[3])
        gpi = g.get("pi")
        if gpi is not None:
            gpi = gpi.reshape(pi.shape)
        return g.get("inputs"), g.get("A"), gpi, g.get("Bfull"), None, None, None


def genepred_run_host(inputs, A, pi, Bfull, codon, end_hints=None, training=False, no_loglik=False,
                      want_post=True, want_alpha=False, want_beta=False):
    """The fused evaluation for HOST inputs (k,B,T,20) (CPU tensor, ideally pinned): batch-sliced H2D -> kernels -> D2H
    pipeline inside the library (hmm_b200_genepred_run_host).  Parameters A, pi, Bfull, codon are CUDA tensors.
    Returns pinned CPU tensors ('post','alpha','beta' (k,B,T,15) as requested, 'loglik' (k,B)), complete on return."""
    lib = _lib.load()
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T, 20), got {tuple(inputs.shape)}")
    if inputs.is_cuda or inputs.dtype != torch.float32:
        raise ValueError("genepred_run_host takes a float32 CPU tensor (use genepred_run for CUDA tensors)")
    inputs = inputs.contiguous()
    k, B, T, _ = inputs.shape
    A = _req(A, "A", (k, 15, 15)); pi = _req(pi.reshape(k, 15), "pi", (k, 15))
    dev = A.device
    Bfull = _req(Bfull, "Bfull", (k, 15, 15)); codon = _codon5(codon)
    if end_hints is not None:
        if end_hints.is_cuda or tuple(end_hints.shape) != (k, B, 2, 15):
            raise ValueError("end_hints must be a CPU tensor of shape (num_models, batch, 2, 15) on the host path")
        end_hints = end_hints.to(torch.float32).contiguous()
    with torch.cuda.device(dev):
        res = {}
        for name, want in (("post", want_post), ("alpha", want_alpha), ("beta", want_beta)):
            if want:
                res[name] = torch.empty((k, B, T, 15), dtype=torch.float32, pin_memory=True)
        res["loglik"] = torch.empty((k, B), dtype=torch.float32, pin_memory=True)
        scratch = _workspace(lib.hmm_b200_genepred_host_scratch_size(k, B, T), dev)
        flags = (TRAINING if training else 0) | (NO_LOGLIK if no_loglik else 0)
        _lib.check(lib.hmm_b200_genepred_run_host(
            _ptr(inputs), k, B, T, _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints), flags,
            _ptr(res.get("post")), _ptr(res.get("alpha")), _ptr(res.get("beta")), _ptr(res["loglik"]),
            _ptr(scratch), scratch.numel(), _stream()))
        torch.cuda.current_stream(dev).synchronize()
    return res


def _shard_common(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints):
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T_local, 20), got {tuple(inputs.shape)}")
    k, B, T, _ = inputs.shape
    if seq_off < 0 or seq_off + T > seq_total:
        raise ValueError(f"shard [{seq_off}, {seq_off + T}) lies outside a sequence of length {seq_total}")
    inputs = _req(inputs, "inputs")
    if halo is None:
        if seq_off != 0 or T != seq_total:
            raise ValueError("a proper shard needs the halo nucleotides of its neighbours (sharded.slice_with_halo)")
    else:
        halo = _req(halo, "halo", (k * B, 4, 5))
    A = _req(A, "A", (k, 15, 15)); pi = _req(pi.reshape(k, 15), "pi", (k, 15))
    Bfull = _req(Bfull, "Bfull", (k, 15, 15)); codon = _codon5(codon)
    if end_hints is not None:
        end_hints = _req(end_hints, "end_hints", (k, B, 2, 15))
    return k, B, T, inputs, halo, A, pi, Bfull, codon, end_hints


def genepred_shard_reduce(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints=None, training=False):
    """Step 1 of the sequence-sharded evaluation (include/hmm_b200.h): this rank's shard (k,B,T_local,20) reduced to
    one stored transfer block per sequence.  Returns (block (k*B,256), workspace) — the workspace holds the shard's
    product trees and must be handed unchanged to genepred_shard_finalize."""
    lib = _lib.load()
    k, B, T, inputs, halo, A, pi, Bfull, codon, end_hints = _shard_common(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints)
    dev = inputs.device
    with torch.cuda.device(dev):
        block = torch.empty((k * B, 256), dtype=torch.float32, device=dev)
        ws = torch.empty(max(lib.hmm_b200_genepred_workspace_size(k, B, T), 256), dtype=torch.uint8, device=dev)
        _lib.check(lib.hmm_b200_genepred_shard_reduce(
            _ptr(inputs), k, B, T, seq_off, seq_total, _ptr(halo), _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints),
            TRAINING if training else 0, _ptr(block), _ptr(ws), ws.numel(), _stream()))
    return block, ws


def genepred_shard_finalize(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, all_blocks, rank, world, workspace,
                            end_hints=None, training=False, no_loglik=False, want_post=True, want_alpha=False,
                            want_beta=False):
    """Step 3: all_blocks (world, k*B, 256) = the all-gathered blocks of step 1.  Returns the requested tensors for
    the local positions and 'loglik' (k,B) of the whole sequence."""
    lib = _lib.load()
    k, B, T, inputs, halo, A, pi, Bfull, codon, end_hints = _shard_common(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints)
    dev = inputs.device
    all_blocks = _req(all_blocks, "all_blocks", (world, k * B, 256))
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    with torch.cuda.device(dev):
        res = {}
        for name, want in (("post", want_post), ("alpha", want_alpha), ("beta", want_beta)):
            if want:
                res[name] = torch.empty((k, B, T, 15), dtype=torch.float32, device=dev)
        res["loglik"] = torch.empty((k, B), dtype=torch.float32, device=dev)
        flags = (TRAINING if training else 0) | (NO_LOGLIK if no_loglik else 0)
        _lib.check(lib.hmm_b200_genepred_shard_finalize(
            _ptr(inputs), k, B, T, seq_off, seq_total, _ptr(halo), _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints),
            flags, _ptr(all_blocks), rank, world, _ptr(res.get("post")), _ptr(res.get("alpha")), _ptr(res.get("beta")),
            _ptr(res["loglik"]), _ptr(workspace), workspace.numel(), _stream()))
    return res


def genepred_shard_reduce_x(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, workspace, step_tag, block_out=None,
                            peer_blocks=None, peer_flags=None, rank=0, world=0, end_hints=None, training=False):
    """Step 1 with the exchange inside the kernel (include/hmm_b200.h): products + shard product; the block goes into
    every rank's table (`peer_blocks` / `peer_flags`: ctypes arrays of device pointers in rank order) and / or to
    `block_out` (k*B,256).  `workspace`: a ZEROED uint8 tensor of hmm_b200_genepred_workspace_size bytes, kept across steps."""
    lib = _lib.load()
    k, B, T, inputs, halo, A, pi, Bfull, codon, end_hints = _shard_common(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints)
    with torch.cuda.device(inputs.device):
        _lib.check(lib.hmm_b200_genepred_shard_reduce_x(
            _ptr(inputs), k, B, T, seq_off, seq_total, _ptr(halo), _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints),
            TRAINING if training else 0, _ptr(block_out), peer_blocks, peer_flags, rank, world, step_tag,
            _ptr(workspace), workspace.numel(), _stream()))


def genepred_shard_finalize_x(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, all_blocks, block_flags, rank, world,
                              step_tag, workspace, end_hints=None, training=False, no_loglik=False, want_post=True,
                              want_alpha=False, want_beta=False, out=None):
    """Step 3 of the same: all_blocks (world, k*B, 256) = this rank's own table half of the step (then block_flags = the
    matching flag half, waited for inside the kernel) or NCCL-gathered blocks (block_flags=None)."""
    lib = _lib.load()
    k, B, T, inputs, halo, A, pi, Bfull, codon, end_hints = _shard_common(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints)
    dev = inputs.device
    with torch.cuda.device(dev):
        res, out = {}, out or {}
        for name, want in (("post", want_post), ("alpha", want_alpha), ("beta", want_beta)):
            if want:
                res[name] = out.get(name) if out.get(name) is not None else torch.empty((k, B, T, 15), dtype=torch.float32, device=dev)
        res["loglik"] = out.get("loglik") if out.get("loglik") is not None else torch.empty((k, B), dtype=torch.float32, device=dev)
        flags = (TRAINING if training else 0) | (NO_LOGLIK if no_loglik else 0)
        _lib.check(lib.hmm_b200_genepred_shard_finalize_x(
            _ptr(inputs), k, B, T, seq_off, seq_total, _ptr(halo), _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints),
            flags, _ptr(all_blocks), _ptr(block_flags), rank, world, step_tag, _ptr(res.get("post")), _ptr(res.get("alpha")),
            _ptr(res.get("beta")), _ptr(res["loglik"]), _ptr(workspace), workspace.numel(), _stream()))
    return res


def genepred_shard_fused_chunk(k, B, T_local, world, device=None):
    """Chunk length of the one-launch shard step for (k, B, T_local) among `world` ranks on `device`, 0 when the shard
    does not fit one launch."""
    lib = _lib.load()
    with torch.cuda.device(device if device is not None else torch.cuda.current_device()):
        return int(lib.hmm_b200_genepred_shard_fused_chunk(k, B, T_local, world))


def genepred_shard_run_x(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, workspace, step_tag, peer_blocks, peer_flags,
                         rank, world, end_hints=None, training=False, no_loglik=False, out=None):
    """The whole sharded step in ONE launch (include/hmm_b200.h, hmm_b200_genepred_shard_run_x): posterior of the local
    positions + log-likelihood of the whole sequences; the blocks travel by peer stores inside the kernel."""
    lib = _lib.load()
    k, B, T, inputs, halo, A, pi, Bfull, codon, end_hints = _shard_common(inputs, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints)
    dev = inputs.device
    with torch.cuda.device(dev):
        out = out or {}
        res = {"post": out.get("post") if out.get("post") is not None else torch.empty((k, B, T, 15), dtype=torch.float32, device=dev),
               "loglik": out.get("loglik") if out.get("loglik") is not None else torch.empty((k, B), dtype=torch.float32, device=dev)}
        flags = (TRAINING if training else 0) | (NO_LOGLIK if no_loglik else 0)
        _lib.check(lib.hmm_b200_genepred_shard_run_x(
            _ptr(inputs), k, B, T, seq_off, seq_total, _ptr(halo), _ptr(A), _ptr(pi), _ptr(Bfull), _ptr(codon), _ptr(end_hints),
            flags, peer_blocks, peer_flags, rank, world, step_tag, _ptr(res["post"]), _ptr(res["loglik"]),
            _ptr(workspace), workspace.numel(), _stream()))
    return res


def genepred_emission(inputs, Bfull, codon, end_hints=None, training=False):
    lib = _lib.load()
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T, 20), got {tuple(inputs.shape)}")
    k, B, T, _ = inputs.shape
    inputs = _req(inputs, "inputs")
    Bfull = _req(Bfull, "Bfull", (k, 15, 15)); codon = _codon5(codon)
    if end_hints is not None:
        end_hints = _req(end_hints, "end_hints", (k, B, 2, 15))
    with torch.cuda.device(inputs.device):
        E = torch.empty((k, B, T, 15), dtype=torch.float32, device=inputs.device)
        _lib.check(lib.hmm_b200_genepred_emission(_ptr(inputs), k, B, T, _ptr(Bfull), _ptr(codon), _ptr(end_hints),
                                                  TRAINING if training else 0, _ptr(E), _stream()))
    return E


def emission_general(inputs, Bfull, codon=None, num_copies=1, end_hints=None, training=False):
    """Emissions of SimpleGenePredHMMEmitter (codon=None: inputs (k,B,T,s)) or GenePredHMMEmitter with any number of
    copies (codon = (2,9,64) codon_probs or the prepared tables: inputs (k,B,T,s+5)); Bfull (k,q,s).  -> (k,B,T,q)."""
    lib = _lib.load()
    if inputs.dim() != 4:
        raise ValueError(f"inputs must have shape (num_models, batch, T, s), got {tuple(inputs.shape)}")
    inputs = _req(inputs, "inputs")
    k, B, T, w = inputs.shape
    has_codon = codon is not None
    s = w - 5 if has_codon else w
    Bfull = _req(Bfull, "Bfull")
    if Bfull.dim() != 3 or Bfull.shape[0] != k or Bfull.shape[2] != s:
        raise ValueError(f"Bfull must have shape ({k}, q, {s}), got {tuple(Bfull.shape)}")
    q = Bfull.shape[1]
    if has_codon:
        codon = _codon5(codon)
        if q != 1 + 14 * num_copies:
            raise ValueError(f"a codon emitter with {num_copies} copies has {1 + 14 * num_copies} states, Bfull has {q}")
    if end_hints is not None:
        end_hints = _req(end_hints, "end_hints", (k, B, 2, q))
    with torch.cuda.device(inputs.device):
        E = torch.empty((k, B, T, q), dtype=torch.float32, device=inputs.device)
        _lib.check(lib.hmm_b200_emission_general(_ptr(inputs), k, B, T, s, q, num_copies, 1 if has_codon else 0, _ptr(Bfull),
                                                 _ptr(codon), _ptr(end_hints), TRAINING if training else 0, _ptr(E), _stream()))
    return E


def dense_run(E, A, pi, want_post=False, want_alpha=False, want_beta=False, no_loglik=False, alpha_normalized=False):
    """Generic dense HMM evaluation from materialised emissions E (k,B,T,q), A (k,q,q), pi (k,q).
    Returns dict with the requested tensors ('post','alpha','beta' each (k,B,T,q)) and 'loglik' (k,B)."""
    lib = _lib.load()
    if not isinstance(E, torch.Tensor) or E.dim() != 4:
        raise ValueError("E must be a (num_models, batch, T, q) tensor")
    k, B, T, q = E.shape
    E = _req(E, "E")
    dev = E.device
    A = _req(A, "A", (k, q, q)); pi = _req(pi.reshape(k, q), "pi", (k, q))
    need_b = want_post or want_beta
    At = A.transpose(1, 2).contiguous() if need_b else None
    with torch.cuda.device(dev):
        res = {"loglik": torch.empty((k, B), dtype=torch.float32, device=dev)}
        alpha = torch.empty((k, B, T, q), dtype=torch.float32, device=dev) if (want_alpha or want_post) else None
        beta = torch.empty((k, B, T, q), dtype=torch.float32, device=dev) if need_b else None
        post = torch.empty((k, B, T, q), dtype=torch.float32, device=dev) if want_post else None
        ws = _workspace(lib.hmm_b200_dense_workspace_size(k, B, T, q), dev)
        flags = (NO_LOGLIK if no_loglik else 0) | (0 if want_alpha else ALPHA_SCRATCH) | (0 if want_beta else BETA_SCRATCH) \
            | (ALPHA_NORMALIZED if alpha_normalized else 0)
        _lib.check(lib.hmm_b200_dense_run(_ptr(E), _ptr(A), _ptr(At), _ptr(pi), k, B, T, q, flags,
                                          _ptr(post), _ptr(alpha), _ptr(beta), _ptr(res["loglik"]), _ptr(ws), ws.numel(), _stream()))
    if want_post:
        res["post"] = post
    if want_alpha:
        res["alpha"] = alpha
    if want_beta:
        res["beta"] = beta
    return res


def dense_run_onehot(symbols, Bt, A, pi, want_post=False, want_alpha=False, want_beta=False, no_loglik=False):
    """dense_run with the one-hot emission contraction fused (include/hmm_b200.h): symbols (k,B,T) uint8, Bt (k, s, q) =
    the emission table transposed (Bt[m][x] = the emissions of every state for symbol x).  No (k,B,T,q) emission tensor."""
    lib = _lib.load()
    if not isinstance(symbols, torch.Tensor) or symbols.dim() != 3 or symbols.dtype != torch.uint8 or not symbols.is_cuda:
        raise ValueError("symbols must be a CUDA uint8 tensor of shape (num_models, batch, T)")
    k, B, T = symbols.shape
    symbols = symbols.contiguous()
    Bt = _req(Bt, "Bt")
    if Bt.dim() != 3 or Bt.shape[0] != k:
        raise ValueError(f"Bt must have shape ({k}, num_symbols, q), got {tuple(Bt.shape)}")
    nsym, q = int(Bt.shape[1]), int(Bt.shape[2])
    dev = symbols.device
    A = _req(A, "A", (k, q, q)); pi = _req(pi.reshape(k, q), "pi", (k, q))
    need_b = want_post or want_beta
    At = A.transpose(1, 2).contiguous() if need_b else None
    with torch.cuda.device(dev):
        res = {"loglik": torch.empty((k, B), dtype=torch.float32, device=dev)}
        alpha = torch.empty((k, B, T, q), dtype=torch.float32, device=dev) if (want_alpha or want_post) else None
        beta = torch.empty((k, B, T, q), dtype=torch.float32, device=dev) if need_b else None
        post = torch.empty((k, B, T, q), dtype=torch.float32, device=dev) if want_post else None
        ws = _workspace(lib.hmm_b200_dense_workspace_size_onehot(k, B, T, q), dev)
        flags = (NO_LOGLIK if no_loglik else 0) | (0 if want_alpha else ALPHA_SCRATCH) | (0 if want_beta else BETA_SCRATCH)
        _lib.check(lib.hmm_b200_dense_run_onehot(_ptr(symbols), _ptr(Bt), nsym, _ptr(A), _ptr(At), _ptr(pi), k, B, T, q, flags,
                                                 _ptr(post), _ptr(alpha), _ptr(beta), _ptr(res["loglik"]), _ptr(ws), ws.numel(),
                                                 _stream()))
    if want_post:
        res["post"] = post
    if want_alpha:
        res["alpha"] = alpha
    if want_beta:
        res["beta"] = beta
    return res


class DenseLoglik(torch.autograd.Function):
    """loglik (k,B) of the generic dense HMM, differentiable w.r.t. E (k,B,T,q), A (k,q,q) and pi (k,q) — training
    profile-HMM models through MsaHmmLayer.forward (the reference differentiates its Python loop, MsaHmmCell.py:73-106).
    Forward = one dense_run keeping log posterior and normalised forward vectors; backward = the expected-count closed
    form: dE = gamma / E, dA = sum_t alpha-hat_{t-1}^T (gamma_t / (alpha-hat_{t-1} A)), dpi = gamma_0 / pi,
    floored factors pass no gradient.  The two contractions over (batch x time) are plain library GEMMs (cuBLAS)."""
    EPS = 1e-16

    @staticmethod
    def forward(ctx, E, A, pi):
        r = dense_run(E, A, pi, want_post=True, want_alpha=True, alpha_normalized=True)
        ctx.save_for_backward(E, A, pi, r["post"], r["alpha"])
        return r["loglik"]

    @staticmethod
    def backward(ctx, g):
        E, A, pi, post, ahat = ctx.saved_tensors
        k, B, T, q = E.shape
        eps = DenseLoglik.EPS
        gam = torch.exp(post) * g.reshape(k, B, 1, 1).to(torch.float32)
        gE = gA = gpi = None
        if ctx.needs_input_grad[0]:
            gE = torch.where(E > eps, gam / E.clamp_min(eps), torch.zeros_like(gam))
        if ctx.needs_input_grad[2]:
            p = pi.reshape(k, 1, q)
            gpi = torch.where(p > eps, gam[:, :, 0] / p.clamp_min(eps), torch.zeros_like(gam[:, :, 0])).sum(1).reshape(pi.shape)
        if ctx.needs_input_grad[1]:
            if T > 1:
                ah = torch.exp(ahat[:, :, :-1]).reshape(k, B * (T - 1), q)
                R = torch.bmm(ah, A)
                gd = torch.where(R > eps, gam[:, :, 1:].reshape(k, B * (T - 1), q) / R.clamp_min(eps), torch.zeros_like(R))
                gA = torch.bmm(ah.transpose(1, 2), gd)
            else:
                gA = torch.zeros_like(A)
        return gE, gA, gpi


def cell_step(emission_probs, scaled, loglik, A, num_models, reverse=False, init=False):
    """One HmmCell.forward call (reference MsaHmmCell.py:73-106).  emission_probs (k*n, q); scaled (k*n, qc*q);
    loglik (k*n, qc); A (k,q,q) already oriented.  Returns (output (k*n, qc*q+qc), new_scaled, new_loglik)."""
    lib = _lib.load()
    A = _req(A, "A")
    q = A.shape[-1]
    E = _req(emission_probs, "emission_probs").reshape(-1, q)
    rows = E.shape[0]
    scaled = _req(scaled, "scaled").reshape(rows, -1)
    qc = scaled.shape[1] // q
    loglik = _req(loglik, "loglik").reshape(rows, qc)
    if rows % num_models or scaled.shape[1] != qc * q or qc not in (1, q):
        raise ValueError(f"inconsistent cell state shapes: E {tuple(E.shape)}, scaled {tuple(scaled.shape)}, q={q}")
    dev = E.device
    with torch.cuda.device(dev):
        out = torch.empty((rows, qc * q + qc), dtype=torch.float32, device=dev)
        ns = torch.empty((rows, qc * q), dtype=torch.float32, device=dev)
        nl = torch.empty((rows, qc), dtype=torch.float32, device=dev)
        _lib.check(lib.hmm_b200_cell_step(_ptr(E), _ptr(scaled), _ptr(loglik), _ptr(A.contiguous()), num_models, rows // num_models,
                                          qc, q, 1 if reverse else 0, 1 if init else 0, _ptr(out), _ptr(ns), _ptr(nl), _stream()))
    return out, ns, nl
