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


class ShardContext:
    """Persistent state of a rank for repeated sequence-sharded evaluations of one shard shape: the zeroed workspace, the
    step counter and -- when the ranks' devices can map each other's memory (torch symmetric memory: CUDA VMM handles
    over NVLink) -- the block / flag tables every rank's kernels store into directly: ONE launch per step when the
    shard's tiles fit the device at once and only the posterior is wanted (hmm_b200_genepred_shard_run_x), else 4 launches;
    no collective call on the data path either way (a rank may take either route, the tables are the same).  Without peer
    mapping the blocks travel by ONE NCCL all-gather between the two halves (5 launches).  `run` must be called by every
    rank of the group, in lockstep."""

    def __init__(self, shard, halo, seq_off, seq_total, A, pi, Bfull, codon, group=None, use_peer_memory=True, one_launch=True):
        import ctypes
        from . import _lib
        if not dist.is_initialized():
            raise RuntimeError("ShardContext needs an initialised torch.distributed process group")
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.halo, self.seq_off, self.seq_total = halo, seq_off, seq_total
        self.params = (A, pi, Bfull, ops._codon5(codon))
        k, B, T, _ = shard.shape
        self.NB, dev = k * B, shard.device
        self.ws = torch.zeros(max(_lib.load().hmm_b200_genepred_workspace_size(k, B, T), 256), dtype=torch.uint8, device=dev)
        self.tag = 0
        self.peer_blocks = self.peer_flags = None
        self.one_launch = False
        n_blk, n_flg = 2 * self.world * self.NB * 256, 2 * self.world * self.NB
        self.n_blk = n_blk
        self.exchange = "NCCL all_gather_into_tensor (1 KB per rank and sequence) between the two kernel pairs"
        self.launches_per_step = 5
        if use_peer_memory and self.world > 1:
            try:
                import torch.distributed._symmetric_memory as symm
                gname = (group or dist.group.WORLD).group_name
                # one allocation: [blocks: 2 x world x NB x 256 fp32][flags: 2 x world x NB u32], zero-filled
                self.table = symm.empty(n_blk + n_flg, dtype=torch.float32, device=dev)
                self.table.zero_()
                hdl = symm.rendezvous(self.table, gname)
                ptrs = [int(p) for p in hdl.buffer_ptrs]
                arr = ctypes.c_ulonglong * self.world
                self.peer_blocks = arr(*ptrs)
                self.peer_flags = arr(*[p + 4 * n_blk for p in ptrs])
                self._hdl = hdl
                torch.cuda.synchronize(dev)
                dist.barrier(group)            # every rank's table is zero before anybody stores into it
                self.exchange = "peer stores over NVLink into symmetric memory, inside the kernels (no collective call)"
                self.launches_per_step = 4
                if one_launch and ops.genepred_shard_fused_chunk(k, B, T, self.world, dev) > 0:
                    self.one_launch = True
                    self.launches_per_step = 1    # posterior-only steps; alpha / beta requests take the 4-launch route
            except Exception as e:             # no peer mapping on this system: the NCCL route
                self.peer_blocks = self.peer_flags = None
                self.exchange += f" [peer memory unavailable: {type(e).__name__}: {e}]"
        if self.peer_blocks is None:
            self.block = torch.empty((self.NB, 256), dtype=torch.float32, device=dev)

    def _run_one_launch(self, shard, A, pi, Bfull, codon, tag, end_hints, training, no_loglik, out):
        return ops.genepred_shard_run_x(shard, self.halo, self.seq_off, self.seq_total, A, pi, Bfull, codon, self.ws, tag,
                                        self.peer_blocks, self.peer_flags, self.rank, self.world, end_hints=end_hints,
                                        training=training, no_loglik=no_loglik, out=out)

    def run(self, shard, end_hints=None, training=False, no_loglik=False, want_post=True, want_alpha=False, want_beta=False, out=None):
        A, pi, Bfull, codon = self.params
        self.tag += 1
        tag = self.tag
        if self.one_launch and want_post and not (want_alpha or want_beta):
            try:
                return self._run_one_launch(shard, A, pi, Bfull, codon, tag, end_hints, training, no_loglik, out)
            except Exception as e:       # the device refused to hold the whole grid at once (MPS share, carve-out): 4 launches
                if "resident at once" not in str(e):
                    raise
                # (nothing was launched and no rank depends on this rank's route: the tables and tags are the same)
                self.one_launch, self.launches_per_step = False, 4
        if self.peer_blocks is not None:
            ops.genepred_shard_reduce_x(shard, self.halo, self.seq_off, self.seq_total, A, pi, Bfull, codon, self.ws, tag,
                                        peer_blocks=self.peer_blocks, peer_flags=self.peer_flags, rank=self.rank, world=self.world,
                                        end_hints=end_hints, training=training)
            half = (tag & 1) * self.world * self.NB
            blocks = self.table[half * 256:(half + self.world * self.NB) * 256].view(self.world, self.NB, 256)
            flags = self.table[self.n_blk + half:self.n_blk + half + self.world * self.NB].view(torch.int32)
        else:
            ops.genepred_shard_reduce_x(shard, self.halo, self.seq_off, self.seq_total, A, pi, Bfull, codon, self.ws, tag,
                                        block_out=self.block, end_hints=end_hints, training=training)
            blocks, flags = all_gather_blocks(self.block, self.group), None
        return ops.genepred_shard_finalize_x(shard, self.halo, self.seq_off, self.seq_total, A, pi, Bfull, codon, blocks, flags,
                                             self.rank, self.world, tag, self.ws, end_hints=end_hints, training=training,
                                             no_loglik=no_loglik, want_post=want_post, want_alpha=want_alpha, want_beta=want_beta,
                                             out=out)


def sequence_sharded_run_local(shards, A, pi, Bfull, codon, **kw):
    """Single-process form (tests, one GPU): `shards` = [(shard, halo, seq_off, seq_total)] for every virtual rank, all
    on this device; the all-gather is a stack.  Same kernels, same order of operations as the multi-rank path."""
    world = len(shards)
    reduced = [ops.genepred_shard_reduce(s, h, off, tot, A, pi, Bfull, codon, end_hints=kw.get("end_hints"),
                                         training=kw.get("training", False)) for (s, h, off, tot) in shards]
    all_blocks = torch.stack([b for b, _ in reduced])
    return [ops.genepred_shard_finalize(s, h, off, tot, A, pi, Bfull, codon, all_blocks, r, world, reduced[r][1], **kw)
            for r, (s, h, off, tot) in enumerate(shards)]
