"""Genome-length sequences sharded over the GPUs of one box (BASELINE config 5).

The reference's own long-sequence device is `parallel_factor`: cut the sequence into chunks, run every chunk for all
q conditioning states, then combine the chunk-end matrices with a TotalProbabilityCell scan
(reference MsaHMMLayer.py:285-319, 384-419; TotalProbabilityCell.py:30-49).  Here the chunks of the first level are
the ranks: every rank holds a contiguous shard, reduces it to ONE 16x16 stored transfer block per sequence, a single
all-gather (1 KB per rank and sequence, NCCL over NVLink) hands every rank all blocks, and each rank folds them
into the boundary vectors entering / leaving its shard and finalises its own positions.  Batch-sharded workloads
(configs 2-4) never come here: independent sequences need no collective.

The 3-mer emission windows reach two nucleotides across a shard edge (reference gene_pred_hmm_emitter.py:245-249):
the caller supplies them as `halo` when it scatters the sequence (slice_with_halo), so the all-gather stays the only
collective on the data path."""
import torch
import torch.distributed as dist

from . import ops


def shard_bounds(seq_total, world):
    """Contiguous, near-equal shards: [(offset, length)] * world; every shard non-empty (needs seq_total >= world)."""
    if world < 1 or seq_total < world:
        raise ValueError(f"cannot cut a sequence of length {seq_total} into {world} non-empty shards")
    base, rem = divmod(seq_total, world)
    out, off = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((off, n))
        off += n
    return out


def slice_with_halo(inputs, rank, world):
    """inputs (k, B, T_total, 20) (host or device): this rank's shard and its halo.

    Returns (shard (k,B,T_local,20) contiguous, halo (k*B,4,5), seq_off, seq_total).  Halo rows = nucleotide flags
    of positions seq_off-2, seq_off-1, seq_off+T_local, seq_off+T_local+1; rows outside the sequence are zero and
    ignored by the kernels (they pad with 1/4 like kmer.py:27-33)."""
    if inputs.dim() != 4 or inputs.shape[-1] != 20:
        raise ValueError(f"inputs must have shape (num_models, batch, T, 20), got {tuple(inputs.shape)}")
    k, B, T, _ = inputs.shape
    off, n = shard_bounds(T, world)[rank]
    shard = inputs[:, :, off:off + n].contiguous()
    halo = torch.zeros((k * B, 4, 5), dtype=inputs.dtype, device=inputs.device)
    nuc = inputs[..., 15:].reshape(k * B, T, 5)
    for slot, t in enumerate((off - 2, off - 1, off + n, off + n + 1)):
        if 0 <= t < T:
            halo[:, slot] = nuc[:, t]
    return shard, halo, off, T


def all_gather_blocks(block, group=None):
    """The one collective of the sharded path: (k*B, 256) per rank -> (world, k*B, 256) on every rank."""
    world = dist.get_world_size(group)
    out = torch.empty((world,) + tuple(block.shape), dtype=block.dtype, device=block.device)
    # concatenated-along-dim-0 output view: accepted by both the NCCL and the gloo backend
    dist.all_gather_into_tensor(out.view((world * block.shape[0],) + tuple(block.shape[1:])), block.contiguous(), group=group)
    return out


def sequence_sharded_run(shard, halo, seq_off, seq_total, A, pi, Bfull, codon, group=None, end_hints=None,
                         training=False, no_loglik=False, want_post=True, want_alpha=False, want_beta=False):
    """Posterior / log alpha / log beta for the local positions of sequences sharded over the ranks of `group`, and
    the log-likelihood (k,B) of the whole sequences (available on every rank)."""
    if not dist.is_initialized():
        raise RuntimeError("sequence_sharded_run needs an initialised torch.distributed process group")
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    block, ws = ops.genepred_shard_reduce(shard, halo, seq_off, seq_total, A, pi, Bfull, codon, end_hints=end_hints,
                                          training=training)
    all_blocks = all_gather_blocks(block, group)
    return ops.genepred_shard_finalize(shard, halo, seq_off, seq_total, A, pi, Bfull, codon, all_blocks, rank, world, ws,
                                       end_hints=end_hints, training=training, no_loglik=no_loglik, want_post=want_post,
                                       want_alpha=want_alpha, want_beta=want_beta)


def sequence_sharded_run_local(shards, A, pi, Bfull, codon, **kw):
    """Single-process form (tests, one GPU): `shards` = [(shard, halo, seq_off, seq_total)] for every virtual rank, all
    on this device; the all-gather is a stack.  Same kernels, same order of operations as the multi-rank path."""
    world = len(shards)
    reduced = [ops.genepred_shard_reduce(s, h, off, tot, A, pi, Bfull, codon, end_hints=kw.get("end_hints"),
                                         training=kw.get("training", False)) for (s, h, off, tot) in shards]
    all_blocks = torch.stack([b for b, _ in reduced])
    return [ops.genepred_shard_finalize(s, h, off, tot, A, pi, Bfull, codon, all_blocks, r, world, reduced[r][1], **kw)
            for r, (s, h, off, tot) in enumerate(shards)]
